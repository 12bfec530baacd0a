"""Parity of the tcgen05 GEMM (echr_gemm_planes) against an fp64 torch matmul of the same planes."""
import pytest
import torch

from echr_b200 import _lib

pytestmark = pytest.mark.gpu


def _split(x, cols_dst):
    lib = _lib.load()
    rows, cols = x.shape
    hi = torch.empty(rows, cols_dst, dtype=torch.bfloat16, device=x.device)
    lo = torch.empty_like(hi)
    _lib.check(lib.echr_split_bf16(x.data_ptr(), rows, cols, x.stride(0), hi.data_ptr(), lo.data_ptr(),
                                   cols_dst, cols_dst, torch.cuda.current_stream().cuda_stream), "split")
    return hi, lo


def _split_f16(x, cols_dst):
    lib = _lib.load()
    rows, cols = x.shape
    out = torch.empty(rows, cols_dst, dtype=torch.float16, device=x.device)
    _lib.check(lib.echr_split_f16(x.data_ptr(), rows, cols, x.stride(0), out.data_ptr(), cols_dst, cols_dst,
                                  torch.cuda.current_stream().cuda_stream), "split_f16")
    return out


def _gemm(a_planes, w_planes, m, n, k, bias, act, block_n, want_bf=False):
    lib = _lib.load()
    dev = a_planes.device
    out = torch.full((m, n), float("nan"), dtype=torch.float32, device=dev)
    oh = ol = None
    if want_bf:
        oh = torch.zeros(m, n, dtype=torch.bfloat16, device=dev)
        ol = torch.zeros(m, n, dtype=torch.bfloat16, device=dev)
    rc = lib.echr_gemm_planes(
        a_planes.data_ptr(), m, k, a_planes.stride(1), a_planes.shape[0], a_planes.stride(0),
        w_planes.data_ptr(), n, w_planes.stride(1), w_planes.shape[0], w_planes.stride(0),
        bias.data_ptr() if bias is not None else None, act, out.data_ptr(), out.stride(0),
        oh.data_ptr() if want_bf else None, ol.data_ptr() if want_bf else None, n, block_n,
        torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "gemm")
    torch.cuda.synchronize()
    return out, oh, ol


@pytest.mark.parametrize("m,n,k,block_n", [
    (128, 128, 64, 128), (128, 128, 512, 128), (256, 512, 1024, 128), (3, 2048, 1536, 128),
    (300, 5748, 1536, 128), (77, 16, 512, 16), (251, 512, 512, 32), (251, 512, 512, 64), (130, 100, 128, 32), (1000, 16, 512, 16), (130, 500, 128, 128),
    (300, 5748, 1536, 256), (4000, 2048, 1024, 256), (129, 300, 64, 256), (20000, 512, 512, 256),
])
@pytest.mark.parametrize("terms", [(1, 1), (2, 1), (2, 2)])
def test_gemm_planes(m, n, k, block_n, terms):
    torch.manual_seed(m * 7 + n * 3 + k)
    dev = torch.device("cuda:0")
    a = torch.randn(m, k, device=dev)
    w = torch.randn(n, k, device=dev) * 0.05
    bias = torch.randn(n, device=dev)
    at, wt = terms
    if at == 1:  # single-plane operands are fp16 (echr_split_f16), A and W alike
        ah, wh = _split_f16(a, k), _split_f16(w, k)
        assert torch.equal(ah, a.half()) and torch.equal(wh, w.half())
        al, wl = ah, wh
    else:
        ah, al = _split(a, k)
        wh, wl = _split(w, k)
    a_pl = torch.stack([ah, al])[:at].contiguous()
    w_pl = torch.stack([wh, wl])[:wt].contiguous()
    out, _, _ = _gemm(a_pl, w_pl, m, n, k, bias, 0, block_n)
    # exact model of what the kernel computes (products of bf16 / fp16 values are exact in fp32; only the
    # accumulation order differs), evaluated in fp64
    A = [ah.double(), al.double()]
    W = [wh.double(), wl.double()]
    ref = A[0] @ W[0].T
    if at == 2:
        ref = ref + A[1] @ W[0].T
    if wt == 2:
        ref = ref + A[0] @ W[1].T
    ref = ref + bias.double()
    err = (out.double() - ref).abs().max().item()
    assert torch.isfinite(out).all()
    assert err < 2e-4 * max(1.0, ref.abs().max().item()), err
    if terms == (2, 2):  # fp32-grade: compare with the unsplit fp32 operands
        full = a.double() @ w.double().T + bias.double()
        err2 = (out.double() - full).abs().max().item()
        assert err2 < 1e-4, err2


def test_gemm_tanh_and_bf16_out():
    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    m, n, k = 200, 512, 512
    a = torch.randn(m, k, device=dev)
    w = torch.randn(n, k, device=dev) * 0.04
    bias = torch.randn(n, device=dev) * 0.1
    ah, al = _split(a, k)
    wh, wl = _split(w, k)
    out, oh, ol = _gemm(torch.stack([ah, al]), torch.stack([wh, wl]), m, n, k, bias, 1, 128, want_bf=True)
    ref = torch.tanh(a.double() @ w.double().T + bias.double())
    assert (out.double() - ref).abs().max().item() < 1e-4
    rec = oh.double() + ol.double()
    assert (rec - out.double()).abs().max().item() < 1e-5


@pytest.mark.parametrize("block_n", [128, 256])
def test_gemm_f16_single_plane_out(block_n):
    """out_hi without out_lo = one fp16 plane (saturating), the operand format of the single-plane precision."""
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.manual_seed(9)
    m, n, k = 1300, 512, 256
    a = torch.randn(m, k, device=dev)
    w = torch.randn(n, k, device=dev) * 0.05
    bias = torch.randn(n, device=dev) * 0.1
    bias[0] = 1e6  # column 0 overflows fp16: must saturate, not become inf
    ah, wh = _split_f16(a, k), _split_f16(w, k)
    out = torch.empty(m, n, device=dev)
    oh = torch.zeros(m, n, dtype=torch.float16, device=dev)
    _lib.check(lib.echr_gemm_planes(ah.data_ptr(), m, k, k, 1, 0, wh.data_ptr(), n, k, 1, 0, bias.data_ptr(), 0,
                                    out.data_ptr(), n, oh.data_ptr(), None, n, block_n,
                                    torch.cuda.current_stream().cuda_stream), "gemm")
    torch.cuda.synchronize()
    ref = ah.double() @ wh.double().T + bias.double()
    assert (out.double() - ref)[:, 1:].abs().max().item() < 2e-4
    assert torch.equal(oh[:, 1:], out[:, 1:].half())
    assert (oh[:, 0] == 65504).all()
