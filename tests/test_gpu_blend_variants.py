"""Every selectable blend-kernel variant against the oracle (images <= 1e-4, mean gradients <= 1e-3, keys and
ranges bit-exact): the knobs are read once per process (csrc/blend.cu, csrc/blend_lq.cu), so each variant runs
__graft_entry__.smoke() in its own interpreter.  Default = lane-per-quad kernels, ring of 8 words, vector REDs."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu

VARIANTS = {
    "quad_of_lanes": dict(ISOGS_FWD_LQ="0", ISOGS_BWD_LQ="0"),                       # DESIGN 3.1 kernels (A/B baseline)
    "quad_of_lanes_ring8": dict(ISOGS_FWD_LQ="0", ISOGS_BWD_LQ="0", ISOGS_NW_FWD="8", ISOGS_NW_BWD="8"),
    "lane_per_quad_ring4": dict(ISOGS_NW_FWD="4", ISOGS_NW_BWD="4"),
    "lane_per_quad_scalar_red": dict(ISOGS_BWD_LQ="2"),
    "lane_per_quad_ids_from_vals": dict(ISOGS_BWD_LQ="3"),
    "lane_per_quad_one_warp_ctas": dict(ISOGS_LQ_WARPS="1"),
}


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_variant_matches_oracle(name):
    env = dict(os.environ)
    for k in ("ISOGS_FWD_LQ", "ISOGS_BWD_LQ", "ISOGS_NW_FWD", "ISOGS_NW_BWD", "ISOGS_LQ_WARPS"):
        env.pop(k, None)
    env.update(VARIANTS[name])
    r = subprocess.run([sys.executable, "-c", "import __graft_entry__ as g; g.smoke()"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, f"{name}: {r.stdout[-2000:]}\n{r.stderr[-4000:]}"
    assert "smoke ok" in r.stdout
