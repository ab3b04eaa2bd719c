"""ipoke_inn_flow_loss_static (include/ipoke_inn.h): the FlowLoss reduction with its scratch in a file-scope __device__
array, for callers that must not allocate device memory.  Called through the C ABI with preallocated buffers and compared
with the existing entry point ipoke_inn_flow_loss_ws on two calls in a row (the shared array keeps the first call's
contents).  Collected last on purpose (the file name).

Tolerance: both entries accumulate in fp64 and round once to float32; they differ only in how many CTAs the fp64
partials are grouped by, so the float32 results agree to 1e-6 relative (the bound tests/test_gpu_parity.py holds the
loss kernel to against the oracle); the gradient outputs are elementwise and must be bit-identical."""
import ctypes

import pytest
import torch

from ipoke_inn_b200 import _capi

MAX_ELEMS = 16777216


def _call_static(L, z, ld, ref, out, dz, dld, nll_w, ld_w, stream):
    B, C = z.shape
    return L.ipoke_inn_flow_loss_static(z.data_ptr(), ld.data_ptr(), ref.data_ptr() if ref is not None else None, B, C,
                                        nll_w, ld_w, out.data_ptr(), dz.data_ptr() if dz is not None else None,
                                        dld.data_ptr() if dld is not None else None, stream)


def test_static_entry_rejects_bad_arguments_without_a_gpu():
    L = _capi.lib()
    assert L.ipoke_inn_flow_loss_static(None, None, None, 4, 8, 1.0, 1.0, None, None, None, None) == -1
    one = ctypes.c_void_p(8)                                   # never dereferenced: the size check comes first
    assert L.ipoke_inn_flow_loss_static(one, one, None, 4097, 4096, 1.0, 1.0, one, None, None, None) == -5
    msg = L.ipoke_inn_last_error_string(-5).decode()
    assert "16777216" in msg and "flow_loss_ws" in msg


@pytest.mark.gpu
def test_static_entry_matches_workspace_entry_twice_in_a_row():
    dev = torch.device("cuda:0")
    L = _capi.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    g = torch.Generator().manual_seed(11)
    ws = torch.zeros(L.ipoke_inn_flow_loss_scratch_bytes() // 8, dtype=torch.float64, device=dev)
    # two different shapes back to back: 40 CTAs then 3 CTAs, so stale partials of the first call lie beyond the second's
    for B, C, nll_w, ld_w in ((20000, 128, 0.7, 1.3), (1500, 100, 1.0, 1.0)):
        z = torch.randn(B, C, generator=g).to(dev)
        ld = torch.randn(B, generator=g).to(dev)
        ref = torch.randn(B, C, generator=g).to(dev)
        out_a, out_b = torch.full((4,), -1.0, device=dev), torch.full((4,), -2.0, device=dev)
        dz_a, dz_b = torch.empty_like(z), torch.empty_like(z)
        dl_a, dl_b = torch.empty_like(ld), torch.empty_like(ld)
        _capi.check(L.ipoke_inn_flow_loss_ws(z.data_ptr(), ld.data_ptr(), ref.data_ptr(), B, C, nll_w, ld_w, out_a.data_ptr(),
                                             dz_a.data_ptr(), dl_a.data_ptr(), ws.data_ptr(), stream), "flow_loss_ws")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        for rep in range(2):                                     # two calls in a row on the shared array
            out_b.fill_(-2.0)
            _capi.check(_call_static(L, z, ld, ref, out_b, dz_b, dl_b, nll_w, ld_w, stream), "flow_loss_static")
            torch.cuda.synchronize()
            a, b = out_a.cpu(), out_b.cpu()
            for k in range(4):
                assert abs(float(b[k]) - float(a[k])) <= 1e-6 * abs(float(a[k])) + 1e-6, (B, rep, k, a, b)
            assert torch.equal(dz_a, dz_b) and torch.equal(dl_a, dl_b)
        assert torch.cuda.memory_allocated(dev) == before        # torch's allocator untouched by the calls
    # nullable outputs and the size limit on a live device
    _capi.check(_call_static(L, z, ld, None, out_b, None, None, 1.0, 1.0, stream), "flow_loss_static")
    torch.cuda.synchronize()
    assert abs(float(out_b[1]) - float(out_a[1])) <= 1e-6 * abs(float(out_a[1]))
    assert L.ipoke_inn_flow_loss_static(z.data_ptr(), ld.data_ptr(), None, MAX_ELEMS // 64 + 1, 64, 1.0, 1.0,
                                        out_b.data_ptr(), None, None, stream) == -5
