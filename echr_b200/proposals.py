"""Proposal selection between the proposal module and the caption path, on the GPU.

Host-side mirror of the reference's three selection functions — same names, arguments, return values and
order — over the CUDA kernels in csrc/kernels_select.cu:
  * gettop1000      eval_utils.py:263-292   top-N score threshold + (n >= k) cells in row-major order
  * gettop1000_nms  eval_utils.py:295-331   greedy 1-D NMS over the (n,k) proposal grid
  * gettopN_nms     eval_utils.py:230-260   sentence-aware NMS used for re-ranking
The reference copies the [T,K] score matrix to the host and walks it in Python double loops; here the matrix
stays in HBM (pass the CUDA tensor the proposal module produced) and only the selected records (at most a few
thousand ints) come back.  `select_topn` / `select_nms` are the batched forms (many videos per launch, device
tensors out) for feeding the path without any round trip.  There is no CPU path: without the CUDA library or
a B200 these functions raise.
"""

import numpy as np
import torch

from . import _lib
from ._lib import check


def _dev(device=None):
    return torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _as_cuda_f32(x, device=None):
    if isinstance(x, torch.Tensor):
        t = x.detach()
        if not t.is_cuda:
            t = t.to(_dev(device))
    else:
        t = torch.as_tensor(np.asarray(x), device=_dev(device))
    return t.to(torch.float32).contiguous()


# ----------------------------------------------------------------------------- batched, device in / device out
def select_topn(scores, mask, clip_offsets, topN=1000, val_score_thres=0.0, cap=None):
    """scores, mask: CUDA fp32 [sum T_v, K] (mask may be None); clip_offsets: int sequence [V+1].
    Returns device tensors (ind, start, end, conf) [V, cap] and count [V] (records of video v in [:count[v]])."""
    lib = _lib.load()
    scores = _as_cuda_f32(scores)
    dev = scores.device
    K = scores.shape[1]
    off = np.asarray(clip_offsets, dtype=np.int32)
    V = off.size - 1
    if cap is None:  # every (n >= k) cell of the longest video: can never overflow
        tmax = int(np.diff(off).max()) if V else 0
        cap = max(1, sum(min(n + 1, K) for n in range(tmax)))
    d_off = torch.as_tensor(off, device=dev)
    mask = _as_cuda_f32(mask, dev) if mask is not None else None
    out_i = [torch.zeros(V, cap, dtype=torch.int32, device=dev) for _ in range(3)]
    conf = torch.zeros(V, cap, dtype=torch.float32, device=dev)
    count = torch.zeros(V, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.echr_select_topn(scores.data_ptr(), mask.data_ptr() if mask is not None else None, d_off.data_ptr(), V, K,
                                   int(topN), float(val_score_thres), out_i[0].data_ptr(), out_i[1].data_ptr(),
                                   out_i[2].data_ptr(), conf.data_ptr(), count.data_ptr(), cap, _stream()),
              "echr_select_topn")
    return out_i[0], out_i[1], out_i[2], conf, count


def select_nms(scores, clip_offsets, overlap=0.8, topN=1000):
    """Batched gettop1000_nms.  Returns device tensors (ind, start, end, score, pick) [V, cap] and count [V]."""
    lib = _lib.load()
    scores = _as_cuda_f32(scores)
    dev = scores.device
    K = scores.shape[1]
    off = np.asarray(clip_offsets, dtype=np.int32)
    V = off.size - 1
    if V and (int(np.diff(off).max()) > 65535 or K > 65535):
        raise ValueError("select_nms packs (n, k) into 16 bits each: videos of more than 65535 clips are not supported")
    per = np.array([lib.echr_select_nms_proposals(int(t), K) for t in np.diff(off)], dtype=np.int64)
    ws_off = np.concatenate([[0], np.cumsum(per)]).astype(np.int64)
    total = int(ws_off[-1])
    cap = max(1, min(int(topN), int(per.max()) if V else 1))
    d_off = torch.as_tensor(off, device=dev)
    d_ws_off = torch.as_tensor(ws_off[:-1].copy(), device=dev)
    ws = torch.empty(max(1, total) * 8, dtype=torch.uint8, device=dev)
    out_i = [torch.zeros(V, cap, dtype=torch.int32, device=dev) for _ in range(4)]
    sc = torch.zeros(V, cap, dtype=torch.float32, device=dev)
    count = torch.zeros(V, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.echr_select_nms(scores.data_ptr(), d_off.data_ptr(), V, K, float(overlap), int(topN), ws.data_ptr(),
                                  d_ws_off.data_ptr(), total, out_i[0].data_ptr(), out_i[1].data_ptr(), out_i[2].data_ptr(),
                                  sc.data_ptr(), out_i[3].data_ptr(), count.data_ptr(), cap, _stream()),
              "echr_select_nms")
    return out_i[0], out_i[1], out_i[2], sc, out_i[3], count


# ----------------------------------------------------------------------------- the reference's call surface
def gettop1000(pred_proposals, tap_masks, cg_gts, duration, featstamp_to_time, val_score_thres=0, topN=1000):
    """eval_utils.py:263-292, same arguments and the same five lists."""
    scores = _as_cuda_f32(pred_proposals)
    nfeats, K = scores.shape
    ind, start, end, conf, count = select_topn(scores, tap_masks, [0, nfeats], topN=topN, val_score_thres=val_score_thres)
    c = int(count[0].item())
    ind, start, end, conf = (t[0, :c].cpu().numpy() for t in (ind, start, end, conf))
    index_select_list = [int(n) for n in ind]
    featstamp_list = [[int(s), int(e)] for s, e in zip(start, end)]
    cg_select_list = [cg_gts[int(n), int(n - s)] for n, s in zip(ind, start)] if len(cg_gts) else []
    timestamp_list = [featstamp_to_time(int(s), int(e), nfeats, duration) for s, e in zip(start, end)]
    confidence = [float(x) for x in conf]
    return index_select_list, featstamp_list, cg_select_list, timestamp_list, confidence


def gettop1000_nms(pred_proposals, tap_masks, cg_gts, duration, featstamp_to_time, overlap=0.8, topN=1000):
    """eval_utils.py:295-331, same arguments and return values (tap_masks is unused there too)."""
    scores = _as_cuda_f32(pred_proposals)
    nfeats, K = scores.shape
    if nfeats < 2:  # the reference indexes an empty proposal array here
        raise IndexError("gettop1000_nms: a video of %d clip(s) has no (n,k) proposal with k < min(n,K)" % nfeats)
    ind, start, end, sc, pick, count = select_nms(scores, [0, nfeats], overlap=overlap, topN=topN)
    c = int(count[0].item())
    ind, start, end, pick = (t[0, :c].cpu().numpy().astype(np.int64) for t in (ind, start, end, pick))
    nms_scores = sc[0, :c].cpu().numpy().astype(np.float64)
    nms_props = np.stack([start, end], 1)
    prop_gts = np.array([])
    if len(cg_gts):
        g = cg_gts.cpu().numpy() if isinstance(cg_gts, torch.Tensor) else np.asarray(cg_gts)
        prop_gts = g[ind, ind - start]
    index_select_list = nms_props[:, 1] - 1
    timestamp_list = [featstamp_to_time(s, e, nfeats, duration) for (s, e) in nms_props]
    return index_select_list, nms_props, prop_gts, timestamp_list, nms_scores


def gettopN_nms(props, prop_scores, sent_score, nms_overlap=0.999, topN=1000, device=None):
    """eval_utils.py:230-260, same arguments and return values (nms_props, nms_scores, pick)."""
    lib = _lib.load()
    props = np.array(props)
    prop_scores = np.array(prop_scores)
    dev = _dev(device)
    n = props.shape[0]
    p64 = torch.as_tensor(np.ascontiguousarray(props[:, :2].T, dtype=np.float64), device=dev)  # [2,n]: t1 row, t2 row
    ps = torch.as_tensor(np.asarray(prop_scores, dtype=np.float64), device=dev)
    ss = torch.as_tensor(np.asarray(sent_score, dtype=np.float64), device=dev)
    off = torch.tensor([0, n], dtype=torch.int32, device=dev)
    cap = max(1, min(int(topN), n if nms_overlap < 1 else int(topN)))
    ws = torch.empty(max(1, n), dtype=torch.uint8, device=dev)
    out = torch.zeros(cap, dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        check(lib.echr_rerank_nms(p64[0].data_ptr(), p64[1].data_ptr(), ps.data_ptr(), ss.data_ptr(), off.data_ptr(), 1,
                                  float(nms_overlap), int(topN), ws.data_ptr(), out.data_ptr(), count.data_ptr(), cap,
                                  _stream()), "echr_rerank_nms")
    pick = [np.int64(i) for i in out[: int(count.item())].cpu().numpy()]
    return props[pick, :], prop_scores[pick], pick
