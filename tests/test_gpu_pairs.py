"""CTA-pair (tcgen05 cta_group::2) contractions on SMALL shapes: the library sends a contraction to pairs only when it has at
least half a wave of 256 x 256 tiles, which none of the parity fixtures reaches -- so the fixtures are run once more in a child
process with PATHS_B200_FORCE_PAIR=1 (read once per process), which routes every eligible contraction (NT with plane operands and
an output wider than 128; stream-K weight gradients whose rows tile by 256) through the pair kernels: partial pair tiles, odd tile
counts, grouped rows, both precision modes."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parity_fixtures_with_every_eligible_contraction_on_cta_pairs():
    env = dict(os.environ, PATHS_B200_FORCE_PAIR="1")
    env.pop("PATHS_B200_NO_PAIR", None)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-q", "-x", "-m", "gpu",
                        "-k", "end2end or level_api or reproducible or fallback", "-p", "no:cacheprovider"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout
