"""GPU unit parity of the stand-alone kernels: fused attention step vs the oracle's `attend`."""
import numpy as np
import pytest
import torch

import helpers
from echr_b200 import _lib
from oracle import echr_oracle as O

pytestmark = pytest.mark.gpu
torch.set_grad_enabled(False)
DEV = "cuda:0"


@pytest.mark.parametrize("staged", [0, 1])
@pytest.mark.parametrize("clip_bf16", [0, 1])
def test_attention_step_matches_oracle(clip_bf16, staged):
    lib = _lib.load()
    sd = helpers.state_dict("bf16w" if clip_bf16 else "fp32")
    a = "lm_model.core.attention."
    g = torch.Generator().manual_seed(3)
    T = 300
    c3d = torch.randn(T, 500, generator=g)
    if clip_bf16:
        c3d = c3d.bfloat16().float()
    lens = [1, 2, 7, 33, 64, 100, 257, 300, 5]
    starts = [0, 299 - 1, 10, 200, 100, 150, 20, 0, 295]
    if staged:  # windows that fit in shared memory: the bulk-copy staged kernel
        lens = [1, 2, 7, 33, 50, 16, 17, 48, 5]
    n = len(lens)
    h = torch.randn(n, 512, generator=g) * 0.3
    lmax = max(lens)
    clip = torch.zeros(n, lmax, 500)
    mask = torch.zeros(n, lmax)
    for i, (s, l) in enumerate(zip(starts, lens)):
        clip[i, :l] = c3d[s:s + l]
        mask[i, :l] = 1
    want = O.attend(sd, h, clip, mask)
    proj = torch.nn.functional.linear(c3d, sd[a + "ctx2att.weight"], sd[a + "ctx2att.bias"])
    q = torch.exp(2 * torch.nn.functional.linear(h, sd[a + "h2att.weight"], sd[a + "h2att.bias"])).to(DEV)
    proj = torch.exp(2 * proj)  # the kernels take e^{2P}, e^{2q} (what the GEMM epilogues write)
    alpha = sd[a + "alpha_net.weight"].reshape(-1).contiguous().to(DEV)
    if clip_bf16:
        clip_d = torch.zeros(T, 512, dtype=torch.bfloat16, device=DEV)
        clip_d[:, :500] = c3d.to(DEV).bfloat16()
        proj_d = proj.to(DEV).bfloat16().contiguous()
        ld = 512
    else:
        clip_d, proj_d, ld = c3d.to(DEV).contiguous(), proj.to(DEV).contiguous(), 500
    rs = torch.tensor(starts, dtype=torch.int32, device=DEV)
    rl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = torch.full((n, 500), float("nan"), device=DEV)
    hi = torch.zeros(n, 512, dtype=torch.bfloat16, device=DEV)
    lo = torch.zeros(n, 512, dtype=torch.bfloat16, device=DEV)
    _lib.check(lib.echr_attention_step(q.data_ptr(), proj_d.data_ptr(), clip_d.data_ptr(), ld, clip_bf16,
                                       alpha.data_ptr(), rs.data_ptr(), rl.data_ptr(), n, out.data_ptr(),
                                       hi.data_ptr(), lo.data_ptr(), 512, max(lens) if staged else 0,
                                       torch.cuda.current_stream().cuda_stream), "attention")
    torch.cuda.synchronize()
    err = (out.cpu() - want).abs().max().item()
    print("attention max err", err)
    assert err < (2e-2 if clip_bf16 else 2e-5), err
    rec = hi.float() + lo.float()
    assert ((rec[:, :500] - out).abs() <= 2e-5 * out.abs() + 1e-7).all()  # hi+lo carries 16 mantissa bits
    assert (rec[:, 500:] == 0).all()


@pytest.mark.parametrize("rows_mult,mean_len", [(1, 0), (40, 0), (40, 30), (40, 8)])
def test_attention_step_f16_matches_oracle(rows_mult, mean_len):
    """The 16-bit bulk-copy staged attention (csrc/attention_f16.cuh: fp16 projection and clip tables, one warp per
    row, online softmax) against the oracle's `attend` on the same fp16-rounded tables: windows of 1 ... 300 clips,
    a few rows and many rows, with 8, 4 and 2 warps per row (chosen from the mean window length the caller states)."""
    lib = _lib.load()
    sd = helpers.state_dict("bf16w")
    a = "lm_model.core.attention."
    g = torch.Generator().manual_seed(5)
    T = 300
    c3d = torch.randn(T, 500, generator=g).half().float()
    lens = [1, 2, 7, 33, 64, 100, 257, 300, 5] * rows_mult
    starts = [0, 299 - 1, 10, 200, 100, 150, 20, 0, 295] * rows_mult
    n = len(lens)
    h = torch.randn(n, 512, generator=g) * 0.3
    P = torch.nn.functional.linear(c3d, sd[a + "ctx2att.weight"], sd[a + "ctx2att.bias"])
    q = torch.nn.functional.linear(h, sd[a + "h2att.weight"], sd[a + "h2att.bias"])
    alpha = sd[a + "alpha_net.weight"].reshape(-1).contiguous()
    # the table format of csrc/attention_f16.cuh: E = e^{2 (P - rowmax)} as fp16, S = e^{2 rowmax} as fp32
    pmax = P.max(1, keepdim=True)[0]
    proj = torch.exp(2 * (P - pmax)).half()
    scale = torch.exp(2 * pmax).reshape(-1).contiguous()
    # oracle on the rounded tables: tanh(x) = 1 - 2 / (e^{2x} + 1) with e^{2(P+q)} = E S e^{2q}
    want = torch.zeros(n, 500)
    for i, (s, l) in enumerate(zip(starts, lens)):
        e2x = proj[s:s + l].double() * scale[s:s + l, None].double() * torch.exp(2 * q[i].double())
        sc = ((1 - 2 / (e2x + 1)) @ alpha.double()).float()
        want[i] = torch.softmax(sc, 0) @ c3d[s:s + l]
    tab_d = torch.zeros(T, 1024, dtype=torch.float16, device=DEV)  # per clip [E (512) | clip (500, zero padded)]
    tab_d[:, :512] = proj.to(DEV)
    tab_d[:, 512:1012] = c3d.to(DEV).half()
    rs = torch.tensor(starts, dtype=torch.int32, device=DEV)
    rl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    out = torch.full((n, 500), float("nan"), device=DEV)
    out16 = torch.full((n, 512), float("nan"), dtype=torch.float16, device=DEV)
    q_d, alpha_d, scale_d = torch.exp(2 * q).to(DEV), alpha.to(DEV), scale.to(DEV)
    _lib.check(lib.echr_attention_step_f16(q_d.data_ptr(), tab_d.data_ptr(), scale_d.data_ptr(),
                                           alpha_d.data_ptr(), rs.data_ptr(), rl.data_ptr(), n, out.data_ptr(),
                                           out16.data_ptr(), 512, mean_len, torch.cuda.current_stream().cuda_stream),
               "attention_f16")  # mean_len picks 8 / 4 / 2 warps per row
    torch.cuda.synchronize()
    err = (out.cpu() - want).abs().max().item()
    print("attention_f16 max err", err)
    assert err < 2e-5, err
    assert ((out16[:, :500].float() - out).abs() <= 1e-3 * out.abs() + 1e-6).all()
    assert (out16[:, 500:] == 0).all()
    # against the reference arithmetic on unrounded inputs: the fp16 tables cost < 2e-3
    clip = torch.zeros(9, 300, 500)
    mask = torch.zeros(9, 300)
    for i, (s, l) in enumerate(zip(starts[:9], lens[:9])):
        clip[i, :l] = c3d[s:s + l]
        mask[i, :l] = 1
    ref = O.attend(sd, h[:9], clip, mask)
    assert (out[:9].cpu() - ref).abs().max().item() < 2e-3
