"""north_star's correctness gate, driven by pytest: the reference's own `test-backend-ops test` harness
(oracle/_ref/test-backend-ops, built unmodified from /root/reference/tests/test-backend-ops.cpp) runs every
case of an op on the plugin (GGML_BACKEND_PATH=libggml-b200.so, device B200_0) and on the reference's ggml CPU
backend and compares them with the harness's own NMSE thresholds (MUL_MAT / MUL_MAT_ID 5e-4:
test-backend-ops.cpp:3718-3728, 3854-3864; MUL_MAT_VEC_FUSION 5e-3: :5546-5548; FLASH_ATTN_EXT 5e-4: :6082-6084).

For every op the test asserts: no FAIL line, `N/N tests passed`, and N >= a floor on the number of cases the
plugin CLAIMS (supports_op) — so a supports_op regression cannot hide as "not supported"."""
import os
import re
import subprocess

import pytest

from conftest import REF_DIR, ROOT

pytestmark = pytest.mark.gpu
PLUGIN = os.path.join(ROOT, "llama-mi50.cpp_b200", "libggml-b200.so")
TBO = os.path.join(REF_DIR, "test-backend-ops")

# op (as printed by the harness: ggml_op_desc or the test's own op_desc) -> fewest cases that must be claimed and pass.
# Floors are the counts measured on B200 (gpurun_out/tbo_<op>.log of the round-2 runs), never raised blindly.
FLOORS = {
    # the hot path
    "MUL_MAT": 244,              # 179 quantized-weight cases + the float-weight (router matrix) cases
    "MUL_MAT_ID": 293,
    "FLASH_ATTN_EXT": 1088,
    "MUL_MAT_VEC_FUSION": 130,
    "MUL_MAT_ID_FUSION": 6,
    # decode glue (row G)
    "RMS_NORM": 17, "RMS_NORM_MUL_ADD": 30, "ROPE": 124, "ROPE_SET_ROWS": 24, "SET_ROWS": 75, "GET_ROWS": 37,
    "ADD": 65, "MUL": 57, "DIV": 57, "SCALE": 4, "CPY": 65, "CONT": 36, "SOFT_MAX": 207,
    "SWIGLU": 8, "SWIGLU_OAI": 8, "GEGLU": 8, "REGLU": 8,
    # MoE routing glue (gpt-oss, config C5)
    "ARGSORT": 74, "TOP_K": 319, "ADD_ID": 36, "TOPK_MOE": 192, "SUM_ROWS": 7, "CLAMP": 2,
}


def run_tbo(op, timeout=1500):
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    r = subprocess.run([TBO, "test", "-b", "B200_0", "-o", op], capture_output=True, text=True, timeout=timeout, env=env)
    out = r.stdout + r.stderr
    try:      # keep the log for profiles/ (scratch directory; absent on a read-only tree)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"tbo_{op}.log"), "w") as f:
            f.write(out)
    except OSError:
        pass
    return r.returncode, out


@pytest.mark.parametrize("op", sorted(FLOORS))
def test_backend_ops(op):
    if not os.path.exists(TBO):
        pytest.skip("oracle/_ref/test-backend-ops not built")
    rc, out = run_tbo(op)
    fails = [l for l in out.splitlines() if "FAIL" in l]
    assert not fails, "\n".join(fails[:20])
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m, out[-3000:]
    n_ok, n_run = int(m.group(1)), int(m.group(2))
    not_supported = out.count("not supported")
    print(f"{op}: {n_ok}/{n_run} passed, {not_supported} not supported")
    assert rc == 0, out[-3000:]
    assert n_ok == n_run, out[-3000:]
    assert n_run >= FLOORS[op], f"{op}: only {n_run} cases claimed by supports_op (floor {FLOORS[op]}); {not_supported} not supported"
