"""CPU restatement (torch fp32 on the host cores) of the ECHR caption-generation hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the reference lines it
follows.  The restatement keeps the reference's *work shape* — one video per call, clip features
zero-padded to the longest event, ctx2att re-applied every decode step, a host check for
termination every step — because it is also the timed CPU baseline (`bench.py --impl reference`).
It is functional (weights come in as a plain state_dict), not a copy of the reference's modules.

Pinned against the reference itself in tests/test_oracle_vs_reference.py (authoring container) and
through tests/golden/*.npz everywhere else.  Beam search is NOT pinned by the reference (it has
none); `beam_search` below is this repo's spec, built on the pinned step function.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

N_HEAD = 16
D_MODEL = 512


# ------------------------------------------------------------------------------------------------
# R1/R2: pairwise temporal position features (models/MA_attention_8_NEW.py:51-79), numpy float64
# ------------------------------------------------------------------------------------------------
def position_matrix(soi):
    """soi int[N,2] half-open [s,e) -> float64 [N,N,2]: (max(|c_i-c_j|/len_i, 1e-3), log(len_j/len_i)).
    The reference casts the length to float32 before dividing (MA:70); centres stay float64."""
    soi = np.asarray(soi)
    s = soi[:, 0:1].astype(np.float64)
    e = soi[:, 1:2].astype(np.float64)
    c = 0.5 * (s + e)
    ln = (soi[:, 1:2] - soi[:, 0:1]).astype(np.float32)
    dc = np.maximum(np.abs((c - c.T) / ln), 1e-3)
    dl = np.log(ln.T / ln)
    return np.stack([dc, dl], axis=2)


def position_embedding(pos, dim=D_MODEL, wave=10000.0):
    """float64 [N,N,2] -> float64 [N,N,dim]; per pair: [sin(dc·f) | cos(dc·f) | sin(dl·f) | cos(dl·f)],
    f_k = 100 / wave^(4k/dim), k < dim/4 (MA:51-64: concatenate on the last axis, then reshape)."""
    k = np.arange(dim // 4, dtype=np.float64)
    div = np.power(wave, (4.0 / dim) * k)
    z = (100.0 * pos)[..., None] / div  # [N,N,2,dim/4]
    emb = np.concatenate([np.sin(z), np.cos(z)], axis=3)  # [N,N,2,dim/2]
    return emb.reshape(pos.shape[0], pos.shape[1], dim)


# ------------------------------------------------------------------------------------------------
# R3-R7: event relation encoder (MA:35-49 and MA:101-177)
# ------------------------------------------------------------------------------------------------
def relation_encode(sd, feats, soi, return_parts=False):
    """feats f32 [N,1012] (mean-pooled c3d | tap_feats[ind]) -> event context f32 [N,512]."""
    p = "fusion_model."
    q = p + "enc_attn."
    n = feats.shape[0]
    pe = torch.from_numpy(position_embedding(position_matrix(soi))).to(torch.float32)  # MA:39-41
    x = F.linear(feats, sd[p + "event_emb.weight"], sd[p + "event_emb.bias"])  # MA:44
    hid = torch.tanh(F.linear(pe.reshape(n * n, D_MODEL), sd[q + "pair_pos_fc1.weight"],
                              sd[q + "pair_pos_fc1.bias"]))  # MA:108-113
    w = F.linear(hid, sd[q + "pair_pos_fc2.weight"], sd[q + "pair_pos_fc2.bias"])  # [N*N,16]
    w = w.reshape(n, n, N_HEAD).permute(0, 2, 1)  # [i,g,j]  (MA:113-116)
    dh = D_MODEL // N_HEAD
    qd = F.linear(x, sd[q + "query_1.weight"], sd[q + "query_1.bias"]).reshape(n, N_HEAD, dh)
    kd = F.linear(x, sd[q + "key_1.weight"], sd[q + "key_1.bias"]).reshape(n, N_HEAD, dh)
    aff = torch.einsum("igd,jgd->igj", qd, kd) * (1.0 / math.sqrt(float(dh)))  # MA:138-143
    prob = torch.softmax(w * aff, dim=2)  # fST0 (MA:148-149), softmax over keys (MA:97,160)
    mixed = torch.einsum("igj,jc->igc", prob, x)  # [N,16,512]  (MA:167-169)
    wo = sd[q + "linear_out_1.weight"].reshape(N_HEAD, dh, D_MODEL)  # grouped 1x1 conv (MA:98,173)
    out = torch.einsum("igc,goc->igo", mixed, wo).reshape(n, D_MODEL) + sd[q + "linear_out_1.bias"]
    if return_parts:
        return out, dict(x=x, w=w, q=qd, k=kd, prob=prob)
    return out


# ------------------------------------------------------------------------------------------------
# A2-A4: context assembly (CaptionGenerator.py:87-167)
# ------------------------------------------------------------------------------------------------
def build_context(sd, tap_feats, c3d_feats, lda_feats, ind, soi):
    """-> video [100], event [N,512], clip [N,Lmax,500] (zero padded), clip_mask [N,Lmax]."""
    pooled = torch.stack([c3d_feats[s:e].mean(0) for s, e in soi], 0)  # CaptionGenerator.py:110-114
    feats = torch.cat([pooled, tap_feats[torch.as_tensor(ind, dtype=torch.long)]], 1)  # :121,:127
    event = relation_encode(sd, feats, soi)  # :129
    lmax = max(e - s for s, e in soi)  # :141
    clip = c3d_feats.new_zeros(len(soi), lmax, c3d_feats.shape[1])
    mask = c3d_feats.new_zeros(len(soi), lmax)
    for i, (s, e) in enumerate(soi):  # :148-151
        clip[i, : e - s] = c3d_feats[s:e]
        mask[i, : e - s] = 1
    return lda_feats, event, clip, mask


# ------------------------------------------------------------------------------------------------
# D1-D4: one decode step (models/OldModel_NEW.py:133-137, 376-401, 801-823)
# ------------------------------------------------------------------------------------------------
def attend(sd, h, clip, mask):
    """Late-fusion attention over an event's clip features, driven by the PREVIOUS hidden state of
    stream 1 (OldModel_NEW.py:805,811).  Softmax over the padded length, then mask and renormalise
    (OldModel_NEW.py:394-397); ctx2att is re-applied to the padded clip tensor every call (:381)."""
    a = "lm_model.core.attention."
    proj = F.linear(clip, sd[a + "ctx2att.weight"], sd[a + "ctx2att.bias"])  # [N,L,512]
    hq = F.linear(h, sd[a + "h2att.weight"], sd[a + "h2att.bias"]).unsqueeze(1)
    score = F.linear(torch.tanh(proj + hq), sd[a + "alpha_net.weight"], sd[a + "alpha_net.bias"])
    wgt = torch.softmax(score.squeeze(2), dim=1) * mask
    wgt = wgt / wgt.sum(1, keepdim=True)
    return torch.bmm(wgt.unsqueeze(1), clip).squeeze(1)  # [N,500]


def lstm_cell(sd, layer, x, h, c):
    """nn.LSTMCell arithmetic, gate order i,f,g,o, both biases (OldModel_NEW.py:780-782)."""
    p = "lm_model.core.layer%d." % layer
    gates = (F.linear(x, sd[p + "weight_ih"], sd[p + "bias_ih"]) +
             F.linear(h, sd[p + "weight_hh"], sd[p + "bias_hh"]))
    i, f, g, o = gates.chunk(4, 1)
    c2 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


def decode_step(sd, it, video, event, clip, mask, state, want_logits=False):
    """token ids int64[N] + state (h[3,N,512], c[3,N,512]) -> (logprobs [N,V+1], new state).
    Three independent LSTM streams: [x|event], [x|attended clip], [x|video]; logits from the
    concatenated hidden states (OldModel_NEW.py:801-823,133-137).  Dropouts are identity in eval."""
    h, c = state
    x = sd["lm_model.embed.weight"][it]
    n = x.shape[0]
    h0, c0 = lstm_cell(sd, 0, torch.cat([x, event], 1), h[0], c[0])
    att = attend(sd, h[1], clip, mask)
    h1, c1 = lstm_cell(sd, 1, torch.cat([x, att], 1), h[1], c[1])
    h2, c2 = lstm_cell(sd, 2, torch.cat([x, video.unsqueeze(0).expand(n, -1)], 1), h[2], c[2])
    logits = F.linear(torch.cat([h0, h1, h2], 1), sd["lm_model.logit.weight"], sd["lm_model.logit.bias"])
    out = logits if want_logits else torch.log_softmax(logits, dim=1)
    return out, (torch.stack([h0, h1, h2]), torch.stack([c0, c1, c2]))


def init_state(n, hidden=512):
    """CG_init_feats_type == '' -> zero state (OldModel_NEW.py:72-78)."""
    return torch.zeros(3, n, hidden), torch.zeros(3, n, hidden)


def position_embedding_f32(soi, rows=None):
    """float32 [len(rows) or N, N, 512] position embedding of the given query rows (default: all).  For large N
    (the reference's default topN = 1000, eval_utils.py:21) the float64 [N,N,512] intermediate of MA:39-41 is
    4 GB, so big cases are evaluated in row tiles; the values are those of `position_embedding`."""
    pos = position_matrix(soi)
    if rows is not None:
        pos = pos[rows]
    return torch.from_numpy(position_embedding(pos)).to(torch.float32)


def relation_encode_tiled(sd, feats, soi, tile=64):
    """`relation_encode` evaluated in tiles of query events: same arithmetic per row (every reduction is over
    the keys j of one query row i), bounded memory.  Used for N up to the reference's topN = 1000."""
    p = "fusion_model."
    q = p + "enc_attn."
    n = feats.shape[0]
    dh = D_MODEL // N_HEAD
    x = F.linear(feats, sd[p + "event_emb.weight"], sd[p + "event_emb.bias"])
    qd = F.linear(x, sd[q + "query_1.weight"], sd[q + "query_1.bias"]).reshape(n, N_HEAD, dh)
    kd = F.linear(x, sd[q + "key_1.weight"], sd[q + "key_1.bias"]).reshape(n, N_HEAD, dh)
    wo = sd[q + "linear_out_1.weight"].reshape(N_HEAD, dh, D_MODEL)
    outs = []
    for r0 in range(0, n, tile):
        rows = np.arange(r0, min(n, r0 + tile))
        pe = position_embedding_f32(soi, rows)  # [r,N,512]
        hid = torch.tanh(F.linear(pe.reshape(-1, D_MODEL), sd[q + "pair_pos_fc1.weight"], sd[q + "pair_pos_fc1.bias"]))
        w = F.linear(hid, sd[q + "pair_pos_fc2.weight"], sd[q + "pair_pos_fc2.bias"]).reshape(len(rows), n, N_HEAD).permute(0, 2, 1)
        aff = torch.einsum("igd,jgd->igj", qd[r0:r0 + len(rows)], kd) * (1.0 / math.sqrt(float(dh)))
        prob = torch.softmax(w * aff, dim=2)
        mixed = torch.einsum("igj,jc->igc", prob, x)
        outs.append(torch.einsum("igc,goc->igo", mixed, wo).reshape(len(rows), D_MODEL) + sd[q + "linear_out_1.bias"])
    return torch.cat(outs, 0)


# ------------------------------------------------------------------------------------------------
# S: greedy sampling loop with the reference's bookkeeping (OldModel_NEW.py:139-187)
# ------------------------------------------------------------------------------------------------
def greedy_decode(sd, video, event, clip, mask, seq_length=30):
    """Returns (seq int64 [N,<=30], logp f32 [N,<=30]) or ([], []) like the reference:
      * step t feeds argmax of step t-1's logprobs (token 0 = BOS at t=0)            (:155-159)
      * the step function runs BEFORE the finished check, 31 evaluations in all      (:171)
      * finished rows are zeroed in `seq` but their max-logprob is still appended    (:175-183)
      * the loop stops (and trims the columns) once every row has produced token 0   (:179)."""
    n = event.shape[0]
    state = init_state(n)
    seq, lps = [], []
    logprobs = None
    unfinished = None
    for t in range(seq_length + 1):
        if t == 0:
            it = torch.zeros(n, dtype=torch.long)
        else:
            best, it = torch.max(logprobs, 1)
        logprobs, state = decode_step(sd, it, video, event, clip, mask, state)
        if t >= 1:
            unfinished = (it > 0) if t == 1 else unfinished & (it > 0)
            if int(unfinished.sum()) == 0:
                break
            seq.append(it * unfinished.to(it.dtype))
            lps.append(best)
    if not seq:
        return [], []
    return torch.stack(seq, 1), torch.stack(lps, 1)


def teacher_forced_logprobs(sd, video, event, clip, mask, seq):
    """OldModel.forward (OldModel_NEW.py:98-130): feed seq[:, i], stop at the first all-zero column
    i >= 1; returns [N, steps, V+1] log-probabilities."""
    n = event.shape[0]
    state = init_state(n)
    outs = []
    for i in range(seq.shape[1] - 1):
        it = seq[:, i]
        if i >= 1 and int(it.sum()) == 0:
            break
        lp, state = decode_step(sd, it, video, event, clip, mask, state)
        outs.append(lp)
    return torch.stack(outs, 1)


def caption_video(sd, tap_feats, c3d_feats, lda_feats, ind, soi, seq_length=30):
    """CaptionGenerator.forward(..., mode='eval') (CaptionGenerator.py:17-44)."""
    video, event, clip, mask = build_context(sd, tap_feats, c3d_feats, lda_feats, ind, soi)
    return greedy_decode(sd, video, event, clip, mask, seq_length)


# ------------------------------------------------------------------------------------------------
# Beam search — THIS REPO'S SPEC (parity unpinned: the reference has no beam search,
# OldModel_NEW.py:144-146).  All arithmetic is the pinned `decode_step`.
# ------------------------------------------------------------------------------------------------
def beam_search(sd, video, event, clip, mask, beam=3, seq_length=30, return_all=False, step_fn=None,
                return_trace=False):
    """Per event keep `beam` slots (decode rows n*beam+b).  Spec:
      * t=0: every slot is fed BOS (token 0) from the zero state; only slot 0 counts as live.
      * each step, every LIVE slot proposes its `beam` best next tokens (log-prob descending, equal
        values -> lower token id); a proposal's score is the slot's cumulative log-prob plus the
        token's log-prob (plain fp32 sum, no length normalisation).
      * the `beam` best proposals (score descending, equal scores -> lower (parent slot, rank)) become
        the new slots 0..beam-1, inheriting the parent's LSTM state and history.
      * a new slot whose token is 0 (EOS) — or any new slot at the last step t = seq_length-1 — is a
        finished hypothesis: it is offered to the event's result (kept if its score is strictly
        greater than the current best; so earlier steps / lower slots win ties) and the slot is dead
        for the next step (dead slots propose nothing and are refilled by the next selection).
      * always runs seq_length steps; result = best finished hypothesis.
    Returns (seq int64 [N,seq_length] zero padded, logp f32 [N,seq_length] zero padded incl. the EOS
    term, score f32 [N]).
    step_fn(it, ev, cl, mk, state) -> (logprobs, state): the step function to drive (default: this module's
    pinned `decode_step`; bench.py passes the reference's own `get_logprobs_state`).
    return_trace: additionally a list with one dict per step — `kept_tok` [N,beam,t+1] / `kept_score` [N,beam]
    (the new slots), `reject` [N] (best score that was NOT kept, over all live x (V+1) proposals; -inf if
    none) — what tests/helpers.check_beam uses to decide whether a differing row is excused by a near tie."""
    n = event.shape[0]
    rows = n * beam
    rep = lambda x: x.repeat_interleave(beam, 0)
    ev, cl, mk = rep(event), rep(clip), rep(mask)
    state = init_state(rows)
    it = torch.zeros(rows, dtype=torch.long)
    live = torch.zeros(n, beam, dtype=torch.bool)
    live[:, 0] = True
    score = torch.zeros(n, beam)
    hist_tok = torch.zeros(n, beam, seq_length, dtype=torch.long)
    hist_lp = torch.zeros(n, beam, seq_length)
    best_score = torch.full((n,), -float("inf"))
    best_tok = torch.zeros(n, seq_length, dtype=torch.long)
    best_lp = torch.zeros(n, seq_length)
    done_all = [[] for _ in range(n)]
    trace = []
    for t in range(seq_length):
        if step_fn is None:
            logprobs, state = decode_step(sd, it, video, ev, cl, mk, state)
        else:
            logprobs, state = step_fn(it, ev, cl, mk, state)
        # per-row top-`beam`: sort is stable on (-value, index) through a two-key trick
        vals, idx = torch.sort(logprobs, dim=1, descending=True, stable=True)
        next_val = vals[:, beam].reshape(n, beam)  # each row's best proposal outside its own top-`beam`
        vals, idx = vals[:, :beam].reshape(n, beam, beam), idx[:, :beam].reshape(n, beam, beam)
        reject = torch.full((n,), -float("inf"))
        parent = torch.zeros(n, beam, dtype=torch.long)
        new_tok = torch.zeros(n, beam, dtype=torch.long)
        new_lp = torch.zeros(n, beam)
        new_score = torch.zeros(n, beam)
        for e in range(n):
            cand = []
            for b in range(beam):
                if not bool(live[e, b]):
                    continue
                for r in range(beam):
                    lp = vals[e, b, r]
                    cand.append((-float(score[e, b] + lp), b, r, int(idx[e, b, r]), float(lp),
                                 float(score[e, b] + lp)))
            cand.sort(key=lambda c: (c[0], c[1], c[2]))
            if return_trace:
                outside = [float(score[e, b] + next_val[e, b]) for b in range(beam) if bool(live[e, b])]
                outside += [c[5] for c in cand[beam:]]
                reject[e] = max(outside) if outside else -float("inf")
            for k in range(beam):
                _, b, r, tok, lp, tot = cand[k]
                parent[e, k], new_tok[e, k], new_lp[e, k], new_score[e, k] = b, tok, lp, tot
        # reorder state and histories by parent
        src = (torch.arange(n).unsqueeze(1) * beam + parent).reshape(-1)
        state = (state[0][:, src], state[1][:, src])
        hist_tok = torch.gather(hist_tok, 1, parent.unsqueeze(2).expand(-1, -1, seq_length)).clone()
        hist_lp = torch.gather(hist_lp, 1, parent.unsqueeze(2).expand(-1, -1, seq_length)).clone()
        hist_tok[:, :, t] = new_tok
        hist_lp[:, :, t] = new_lp
        score = new_score
        if return_trace:
            trace.append(dict(kept_tok=hist_tok[:, :, :t + 1].clone(), kept_score=score.clone(), reject=reject))
        fin = (new_tok == 0) | (t == seq_length - 1)
        for e in range(n):
            for k in range(beam):
                if bool(fin[e, k]):
                    if return_all:
                        done_all[e].append((float(score[e, k]), hist_tok[e, k].clone(), hist_lp[e, k].clone()))
                    if float(score[e, k]) > float(best_score[e]):
                        best_score[e] = score[e, k]
                        best_tok[e] = hist_tok[e, k]
                        best_lp[e] = hist_lp[e, k]
        live = ~fin
        it = new_tok.reshape(-1)
    out = (best_tok, best_lp, best_score)
    if return_all:
        out = out + (done_all,)
    if return_trace:
        out = out + (trace,)
    return out


def caption_video_beam(sd, tap_feats, c3d_feats, lda_feats, ind, soi, beam=3, seq_length=30):
    video, event, clip, mask = build_context(sd, tap_feats, c3d_feats, lda_feats, ind, soi)
    return beam_search(sd, video, event, clip, mask, beam, seq_length)


def replay_beam_trace(sd, video, event, clip, mask, parent, tok, beam, seq_length=30):
    """Replays a beam search whose per-step decisions (parent slot, token of every new slot: int [S,N,beam]) were
    made elsewhere — the CUDA path, echr_get_beam_trace — with THIS module's arithmetic, and measures how far each
    decision is from the one the spec (`beam_search`) would have made on the same beam state:
      * `regret` [S,N]: (score of the beam-th best proposal of the step) - (score of the worst proposal the trace
        kept), clipped at 0.  0 = the trace kept a best-`beam` set; a value below the near-tie threshold means the
        two differ by a near tie; anything larger is a wrong decision.
      * `invalid`: steps where the trace chose a dead parent or the same proposal twice (must be empty).
      * the finished hypotheses along the trace (per event: list of (score, tokens[S], logp[S])) and, from them,
        the best one per event by the spec's rule (strictly greater score; earlier step / lower slot wins ties).
    Returns dict(regret, invalid, best_tok [N,S], best_lp [N,S], best_score [N], finished)."""
    n = event.shape[0]
    parent = torch.as_tensor(np.asarray(parent)).long()
    tok = torch.as_tensor(np.asarray(tok)).long()
    rep = lambda x: x.repeat_interleave(beam, 0)
    ev, cl, mk = rep(event), rep(clip), rep(mask)
    state = init_state(n * beam)
    it = torch.zeros(n * beam, dtype=torch.long)
    live = torch.zeros(n, beam, dtype=torch.bool)
    live[:, 0] = True
    score = torch.zeros(n, beam)
    hist_tok = torch.zeros(n, beam, seq_length, dtype=torch.long)
    hist_lp = torch.zeros(n, beam, seq_length)
    regret = torch.zeros(seq_length, n)
    invalid = []
    finished = [[] for _ in range(n)]
    best_score = torch.full((n,), -float("inf"))
    best_tok = torch.zeros(n, seq_length, dtype=torch.long)
    best_lp = torch.zeros(n, seq_length)
    for t in range(seq_length):
        logprobs, state = decode_step(sd, it, video, ev, cl, mk, state)
        lp3 = logprobs.reshape(n, beam, -1)
        total = score.unsqueeze(2) + lp3  # [n, beam, V+1] proposal scores
        total = torch.where(live.unsqueeze(2), total, torch.full_like(total, -float("inf")))
        kth = torch.topk(total.reshape(n, -1), beam, dim=1)[0][:, beam - 1]  # beam-th best proposal
        p_t, k_t = parent[t], tok[t]  # [n, beam]
        chosen = torch.gather(total.reshape(n, -1), 1, p_t * lp3.shape[2] + k_t)  # [n, beam]
        chosen_lp = torch.gather(lp3.reshape(n, -1), 1, p_t * lp3.shape[2] + k_t)
        any_live = live.any(1)
        for e in range(n):
            if not bool(any_live[e]):
                continue  # every slot finished: nothing to choose (only possible after the last step)
            pairs = set()
            for k in range(beam):
                key = (int(p_t[e, k]), int(k_t[e, k]))
                if not bool(live[e, key[0]]) or key in pairs:
                    invalid.append((t, e, k))
                pairs.add(key)
        regret[t] = torch.where(any_live, (kth - chosen.min(1)[0]).clamp(min=0), torch.zeros(n))
        src = (torch.arange(n).unsqueeze(1) * beam + p_t).reshape(-1)
        state = (state[0][:, src], state[1][:, src])
        hist_tok = torch.gather(hist_tok, 1, p_t.unsqueeze(2).expand(-1, -1, seq_length)).clone()
        hist_lp = torch.gather(hist_lp, 1, p_t.unsqueeze(2).expand(-1, -1, seq_length)).clone()
        hist_tok[:, :, t] = k_t
        hist_lp[:, :, t] = chosen_lp
        score = chosen
        fin = ((k_t == 0) | (t == seq_length - 1)) & any_live.unsqueeze(1)
        for e in range(n):
            for k in range(beam):
                if bool(fin[e, k]) and float(score[e, k]) > -float("inf"):
                    finished[e].append((float(score[e, k]), hist_tok[e, k].clone(), hist_lp[e, k].clone()))
                    if float(score[e, k]) > float(best_score[e]):
                        best_score[e] = score[e, k]
                        best_tok[e] = hist_tok[e, k]
                        best_lp[e] = hist_lp[e, k]
        live = ~fin & any_live.unsqueeze(1)
        it = k_t.reshape(-1)
    return dict(regret=regret, invalid=invalid, best_tok=best_tok, best_lp=best_lp, best_score=best_score, finished=finished)
