"""dlp_attn_paged_static: the split-KV attention entry point for hosts that cannot allocate device memory.

Who that is: a server whose allocator is frozen after it divided HBM into weights and KV pages, a plug-in under a
"no allocation after init" rule, code inside a stream capture (cudaMalloc is illegal there).  For them the split-KV
partials live in a fixed 4 MiB array inside the library instead of a caller-provided workspace.

The test drives both entry points through the raw C ABI (ctypes on libdlp_b200.so, no Python wrapper in between) and
requires the static variant to reproduce `dlp_attn_paged` - twice in a row with different inputs, so that leftovers of
the first call in the shared array would show up in the second.  It also checks that an over-sized request is refused
with a message instead of running.  This file sorts last on purpose: a run that stops at the first failure has by then
executed every other test.
"""
import ctypes as C
import math

import pytest
import torch

from distributed_llm_pipeline_b200.models.reference import nmse
from distributed_llm_pipeline_b200.ops import lib as dlp_lib

pytestmark = pytest.mark.gpu
SCRATCH_FLOATS = 1 << 20  # DLP_ATTN_STATIC_SCRATCH_FLOATS in include/dlp_b200.h


def _abi():
    lib = C.CDLL(dlp_lib.LIB_PATH)
    vp, i32 = C.c_void_p, C.c_int
    lib.dlp_attn_paged.restype = i32
    lib.dlp_attn_paged.argtypes = [vp] * 9 + [i32] * 7 + [vp]
    lib.dlp_attn_paged_static.restype = i32
    lib.dlp_attn_paged_static.argtypes = [vp] * 7 + [i32] * 7 + [vp]
    lib.dlp_last_error.restype = C.c_char_p
    return lib


def _case(seed, lens, Hq, Hkv, D, page, max_pages):
    g = torch.Generator(device="cuda").manual_seed(seed)
    nq = len(lens)
    n_pages = nq * max_pages + 3
    kc = torch.randn(n_pages, Hkv, page, D, device="cuda", generator=g).half()
    vc = torch.randn(n_pages, Hkv, page, D, device="cuda", generator=g).half()
    bt = torch.randperm(n_pages)[: nq * max_pages].view(nq, max_pages).to(torch.int32).cuda()
    kv_len = torch.tensor(lens, dtype=torch.int32, device="cuda")
    q = (torch.randn(nq, Hq * D, device="cuda", generator=g) / math.sqrt(D)).half()
    return q, kc, vc, bt, kv_len


def test_static_scratch_matches_workspace_entry_point_twice():
    lib = _abi()
    Hq, Hkv, D, page, max_pages, n_splits = 8, 2, 128, 16, 48, 4
    stream = torch.cuda.current_stream().cuda_stream
    # two different problems back to back: the second must not see the first one's partials
    for seed, lens in ((1, [1, 17, 300, 700]), (2, [650, 3, 264])):   # contexts long enough that the split-KV scratch is really used
        q, kc, vc, bt, kv_len = _case(seed, lens, Hq, Hkv, D, page, max_pages)
        nq = len(lens)
        want = torch.empty(nq, Hq * D, dtype=torch.float16, device="cuda")
        po = torch.empty(nq * Hq * n_splits * D, dtype=torch.float32, device="cuda")
        pm = torch.empty(nq * Hq * n_splits * 2, dtype=torch.float32, device="cuda")
        rc = lib.dlp_attn_paged(q.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), None, kv_len.data_ptr(),
                                want.data_ptr(), po.data_ptr(), pm.data_ptr(), nq, Hq, Hkv, D, max_pages, page, n_splits, stream)
        assert rc == 0, lib.dlp_last_error()
        got = torch.full_like(want, float("nan"))
        rc = lib.dlp_attn_paged_static(q.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), None, kv_len.data_ptr(),
                                       got.data_ptr(), nq, Hq, Hkv, D, max_pages, page, n_splits, stream)
        assert rc == 0, lib.dlp_last_error()
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert nmse(want, got) < 5e-5


def test_static_scratch_rejects_oversized_request():
    lib = _abi()
    Hq, Hkv, D, page, max_pages = 8, 2, 128, 16, 4
    q, kc, vc, bt, kv_len = _case(3, [5, 9], Hq, Hkv, D, page, max_pages)
    out = torch.zeros(2, Hq * D, dtype=torch.float16, device="cuda")
    n_splits = SCRATCH_FLOATS // (2 * Hq * (D + 2)) + 1   # one split too many for the built-in array
    rc = lib.dlp_attn_paged_static(q.data_ptr(), kc.data_ptr(), vc.data_ptr(), bt.data_ptr(), None, kv_len.data_ptr(),
                                   out.data_ptr(), 2, Hq, Hkv, D, max_pages, page, n_splits,
                                   torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    assert b"exceeds the built-in scratch" in lib.dlp_last_error()
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0   # nothing was launched
