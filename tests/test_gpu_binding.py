"""The compiled reference-side binding (-m gpu): integration/_build/lara_gpu is the REFERENCE's own front end (its
SeqAn argument parser, RNA readers and writers, compiled unmodified from the reference tree by integration/build.sh)
with lara::solve bound to liblara_b200.so through the C ABI (integration/subgradient_solver_gpu.hpp).  Running it
on the committed fixtures must give the bytes the unmodified reference binaries wrote: the library is a drop-in at
src/lara.cpp:57, and the reference's parsers in front of the GPU agree with our own host readers."""
import os
import subprocess
import tempfile

import pytest

from tests import util

pytestmark = pytest.mark.gpu

BINARY = os.path.join(util.ROOT, "integration", "_build", "lara_gpu")


@pytest.mark.parametrize("case", util.golden_cases(), ids=lambda c: c["input"] + "-" + c["tag"])
def test_reference_front_end_with_gpu_solver(built, case):
    if not os.path.exists(BINARY):
        pytest.fail("integration/_build/lara_gpu is missing: it is built in the build container by "
                    "__graft_entry__.build() (integration/build.sh needs the reference sources) and travels with the snapshot")
    files = [os.path.join(util.GOLDEN, f) for f in case["files"]]
    argv = []
    if case["format"] == "dp":
        for f in files:
            argv += ["-d", f]
    else:
        argv += ["-i", files[0]]
    if case.get("ref_file"):
        argv += ["-r", os.path.join(util.GOLDEN, case["ref_file"])]
    options = [os.path.join(util.GOLDEN, o[len("@golden/"):]) if o.startswith("@golden/") else o for o in case["options"]]
    out = os.path.join(tempfile.mkdtemp(prefix="lara_bind_"), "out.txt")
    env = dict(os.environ, LARA_B200_STEP_MODE="1" if case["build"] == "simd" else "0")
    cp = subprocess.run([BINARY] + argv + options + ["-j", "1", "-w", out], env=env, stdout=subprocess.PIPE,
                        stderr=subprocess.PIPE, timeout=600)
    assert cp.returncode == 0, cp.stderr.decode()[-800:]
    with open(out, "rb") as f:
        got = f.read()
    with open(os.path.join(util.GOLDEN, case["output"]), "rb") as f:
        want = f.read()
    assert got == want
