"""The CUDA kernels against the golden vectors generated from the reference itself (tests/golden/make_golden.py ->
quant_golden.npz), through the C ABI: q8_1 activation blocks and q8_0 KV rows bit-exact, dequantising GET_ROWS equal up to
fma contraction."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14, "mxfp4": 39}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "quant_golden.npz"))


def test_gpu_q8_1_blocks_match_golden(gold):
    import torch
    from llama_mi50_b200 import capi
    x = gold["x"]
    got = capi.quantize_q8_1(torch.from_numpy(x).cuda(), x.shape[1]).cpu().numpy().reshape(x.shape[0], -1)
    assert np.array_equal(got, gold["q8_1_bytes"])


def test_gpu_set_rows_q8_0_matches_golden(gold):
    import torch
    from llama_mi50_b200 import capi
    x = gold["x"]
    n, k = x.shape
    idx = np.arange(n, dtype=np.int64)[::-1].copy()
    dst = torch.zeros(n * (k // 32) * 34, dtype=torch.uint8, device="cuda")
    capi.set_rows(torch.from_numpy(x).cuda(), torch.from_numpy(idx).cuda(), dst, capi.Q8_0, [k, n, 1, 1])
    got = dst.cpu().numpy().reshape(n, -1)
    for i, r in enumerate(idx):
        assert np.array_equal(got[r], gold["q8_0_bytes"][i]), f"row {i}"


@pytest.mark.parametrize("name", list(TYPES))
def test_gpu_get_rows_matches_golden_dequant(gold, name):
    import torch
    from llama_mi50_b200 import capi
    q, d = gold[f"{name}_bytes"], gold[f"{name}_deq"]
    n, k = d.shape
    ids = np.arange(n, dtype=np.int32)
    got = capi.get_rows(torch.from_numpy(np.ascontiguousarray(q).reshape(-1)).cuda(), TYPES[name], [k, n, 1, 1], torch.from_numpy(ids).cuda()).cpu().numpy()
    for r in range(n):
        scale = np.abs(d[r]).max() or 1.0
        assert np.abs(got[r] - d[r]).max() <= 1e-6 * scale, f"row {r}"   # fma contraction only
