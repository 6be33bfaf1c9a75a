"""CPU tests of the thin host side (include/lara_host.h): option table, input readers, flat layout, writers.
Mirrors the behaviour of src/parameters.hpp, src/io.hpp of the reference."""
import os

import numpy as np
import pytest

from tests import util


def test_option_table_defaults_and_conversions(built):
    import lara_b200 as lb
    p = lb.Parameters(["-i", "x.dbn", "-n", "50", "-a", "7", "-x", "-1.5", "-y", "-5", "-l", "100", "900", "-o", "pairs",
                       "-m", "3", "-u", "25", "-b", "0.5", "-c", "2.0", "-f", "1.25", "-j", "2"])
    assert p.status == lb.Parameters.CONTINUE
    assert (p.num_iterations, p.max_nondecreasing, p.matching) == (50, 7, 3)
    assert p.gap_extend == int(-1.5 * 8192) and p.gap_open == int(-5 * 8192)      # src/parameters.hpp:303-304
    assert (p.lib_score_min, p.lib_score_max, p.lib_score_linear) == (100, 900, 1)
    assert p.out_format == "pairs" and p.threads == 2
    assert abs(p.suboptimal_diff - 25) < 1e-6 and abs(p.balance - 0.5) < 1e-6 and abs(p.step_size_factor - 1.25) < 1e-6


def test_bad_options_are_errors(built):
    import lara_b200 as lb
    assert lb.Parameters(["-o", "nonsense", "-i", "x"]).status == lb.Parameters.EXIT_ERROR
    assert lb.Parameters(["--no-such-option"]).status == lb.Parameters.EXIT_ERROR
    assert lb.Parameters(["-n"]).status == lb.Parameters.EXIT_ERROR


def test_read_ebpseq_truncates_weights_and_builds_symmetric_csr(built):
    import lara_b200 as lb
    case = [c for c in util.golden_cases() if c["input"] == "ebp7x60"][0]
    path = os.path.join(util.GOLDEN, case["files"][0])
    s = lb.InputStorage.from_files(in_file=path)
    assert len(s) == 7
    seq_cat, seq_off, bp_partner, bp_weight, bp_off = s.flat()
    assert seq_cat.max() <= 4 and seq_off[0] == 0 and len(bp_off) == seq_off[-1] + 1
    assert bp_weight.min() >= 1 and bp_weight.max() <= 5          # int-truncated, src/lagrange.hpp:135
    # undirected graph: (i,j,w) present <=> (j,i,w) present
    for k in range(len(s)):
        base = seq_off[k]
        L = s.length(k)
        edges = set()
        for i in range(L):
            for x in range(bp_off[base + i], bp_off[base + i + 1]):
                edges.add((i, int(bp_partner[x]), int(bp_weight[x])))
        assert all((j, i, w) in edges for (i, j, w) in edges)


def test_fixed_structure_weight_is_ten(built):
    """computeStructure (src/io.hpp:290-296): a fixed structure counts as probability 1.0 scaled by 10."""
    import lara_b200 as lb
    case = [c for c in util.golden_cases() if c["input"] == "dbn8x70"][0]
    s = lb.InputStorage.from_files(in_file=os.path.join(util.GOLDEN, case["files"][0]))
    w = s.flat()[3]
    assert len(w) > 0 and set(np.unique(w)) == {10}


def test_dotplot_reader(built):
    import lara_b200 as lb
    case = [c for c in util.golden_cases() if c["input"] == "dp4x90"][0]
    s = lb.InputStorage.from_files(dotplots=[os.path.join(util.GOLDEN, f) for f in case["files"]])
    assert len(s) == 4
    w = s.flat()[3]
    assert w.min() >= 0 and w.max() <= 5                          # ln(p / 0.003) truncated, src/io.hpp:257-258


def test_input_errors(built, tmp_path):
    import lara_b200 as lb
    with pytest.raises(lb.LaraError):
        lb.InputStorage.from_files(in_file=str(tmp_path / "missing.ebpseq"))
    one = tmp_path / "one.dbn"
    one.write_text(">a\nACGU\n....\n")
    with pytest.raises(lb.LaraError):                             # "at least two sequences", src/io.hpp:115-119
        lb.InputStorage.from_files(in_file=str(one))
    nostruct = tmp_path / "x.fasta"
    nostruct.write_text(">a\nACGU\n>b\nACGG\n")
    with pytest.raises(lb.LaraError):                             # needs ViennaRNA in the reference: out of scope
        lb.InputStorage.from_files(in_file=str(nostruct))


def test_add_record_equals_file(built):
    import lara_b200 as lb
    s = lb.InputStorage()
    s.add_record("a", "ACGUACGUAC", [(0, 9, 5.7), (1, 8, 0.9)], is_bpp=True)
    s.add_record("b", "ACGUNCGU", [(0, 7, 1.0)], is_bpp=False)
    s.finish()
    seq_cat, seq_off, bp_partner, bp_weight, bp_off = s.flat()
    assert list(seq_off) == [0, 10, 18]
    assert list(seq_cat[10:18]) == [0, 1, 2, 3, 4, 1, 2, 3]
    # 5.7 -> 5, 0.9 -> 0 (kept as an edge with weight 0), fixed structure -> 10
    assert sorted(bp_weight.tolist()) == [0, 0, 5, 5, 10, 10]
