"""GPU parity against the reference's own outputs (-m gpu): the CUDA path, driven through the C ABI, must write
byte-identical T-Coffee library / pairs / fasta files to the ones the unmodified reference binaries produced
for the committed fixtures (tests/golden/manifest.json), in both step-size arithmetics."""
import os

import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", util.golden_cases(), ids=lambda c: c["input"] + "-" + c["tag"])
def test_gpu_reproduces_reference_output(built, case):
    import lara_b200 as lb
    p, store = util.golden_setup(case)
    ctx = lb.Context(p)
    ctx.set_sequences(store.flat())
    res = ctx.solve_all_pairs()
    got = util.write_results(store, res, p.out_format)
    st = ctx.stats()
    ctx.free(res)
    ctx.close()
    with open(os.path.join(util.GOLDEN, case["output"]), "rb") as f:
        want = f.read()
    assert got == want
    assert st.kernel_launches > 0 and st.dp_cells > 0
