"""fp32 host restatement of every native kernel, on the same tables and buffers.

Two jobs: (1) it is what the ops run on a CPU-only machine (the CPU test-suite, gloo
multi-process tests, ``mpirun -np 3``-style plumbing runs of BASELINE config 1); (2) it is the
oracle the GPU numerics tests compare the CUDA kernels against, entry point by entry point.
Itself it is validated against ``torch.nn`` + ``torch.optim.Adam`` + ``StepLR`` (custom mode)
and ``sklearn.neural_network.MLPClassifier`` (sklearn mode) in tests/test_reference_ops.py.

All functions take CPU torch tensors (viewed as NumPy, zero-copy) and the NumPy structured
tables of ops/_native.py; argument order mirrors include/fedb200.h.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch

from ..models.init import philox4x32
from ._native import FEDSTATE_DTYPE, LOSS_SLOTS, MAX_CLASSES, TILE_ROWS

F32 = np.float32


def _np(t: Optional[torch.Tensor]):
    return None if t is None else t.detach().numpy()


def _layout(g):
    L = int(g["n_layers"])
    dims = [int(d) for d in g["dims"][: L + 1]]
    w_off, b_off, ld, off = [], [], [], 0
    for l in range(L):
        ld.append((dims[l] + 3) // 4 * 4)
        w_off.append(off)
        off += dims[l + 1] * ld[l]
        b_off.append(off)
        off += (dims[l + 1] + 3) // 4 * 4
    return L, dims, w_off, b_off, ld, off


def _act(z, act):
    if act == 0:
        return np.maximum(z, F32(0))
    if act == 1:
        return (F32(1) / (F32(1) + np.exp(-z))).astype(F32)
    if act == 2:
        return np.tanh(z).astype(F32)
    return z


def _act_bwd(a, act):
    if act == 0:
        return (a > 0).astype(F32)
    if act == 1:
        return a * (F32(1) - a)
    if act == 2:
        return F32(1) - a * a
    return np.ones_like(a)


def _forward(g, slab, Xb):
    """Returns list of activations a[0..L] (a[L] = raw output z)."""
    L, dims, w_off, b_off, ld, _ = _layout(g)
    a = [Xb[:, : ld[0]]]
    for l in range(L):
        W = slab[w_off[l]: w_off[l] + dims[l + 1] * ld[l]].reshape(dims[l + 1], ld[l])
        b = slab[b_off[l]: b_off[l] + dims[l + 1]]
        x = a[-1]
        if x.shape[1] < ld[l]:
            x = np.pad(x, ((0, 0), (0, ld[l] - x.shape[1])))
        z = (x @ W.T + b).astype(F32)
        a.append(_act(z, int(g["act"])) if l + 1 < L else z)
    return a


def _batch_bounds(g, batch):
    n = int(g["n_rows"])
    bs = min(int(g["batch_size"]), n) if int(g["batch_size"]) > 0 else n
    b0 = batch * bs
    if b0 >= n:
        return None
    return b0, min(bs, n - b0)


# ---------------------------------------------------------------------------------------
def init_params(groups, max_slab, params, streams, kind, seed):
    p = _np(params)
    for gi, g in enumerate(groups):
        sid = int(streams[gi])
        if sid < 0:
            continue
        L, dims, w_off, b_off, ld, end = _layout(g)
        P = int(g["slab_size"])
        q = np.arange(P // 4, dtype=np.uint32)
        words = philox4x32(q, np.uint32(sid), 0, 0, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        x = np.stack(words, axis=1).reshape(-1)
        u = (x >> np.uint32(8)).astype(F32) * F32(2.0 ** -24)
        bound = np.zeros(P, dtype=F32)
        for l in range(L):
            fi, fo = dims[l], dims[l + 1]
            if kind == 1:
                s = F32(1.0 / math.sqrt(fi))
            elif kind == 2:
                s = F32(math.sqrt((2.0 if int(g["act"]) == 1 else 6.0) / (fi + fo)))
            else:
                s = F32(0)
            bound[w_off[l]: w_off[l] + fo * ld[l]].reshape(fo, ld[l])[:, :fi] = s
            bound[b_off[l]: b_off[l] + fo] = s
        off = int(g["slab_off"])
        p[off: off + P] = (F32(2) * u - F32(1)) * bound


def round_begin(feds, fed_state, conf):
    st = _np(fed_state).view(FEDSTATE_DTYPE)
    for f in range(len(feds)):
        if st[f]["stop_pending"]:
            st[f]["stop"] = 1
        if not st[f]["stop"]:
            st[f]["round"] += 1
    _np(conf)[...] = 0


def _feistel_perm(n, ka, kb):
    bits = 1
    while (1 << bits) < n:
        bits += 1
    half = (bits + 1) >> 1
    mask = np.uint32((1 << half) - 1)
    x = np.arange(n, dtype=np.uint32)
    out = np.empty(n, dtype=np.uint32)
    pending = np.arange(n)
    while len(pending):
        l, r = x >> np.uint32(half), x & mask
        for rnd in range(4):
            c0, _, _, _ = philox4x32(r, np.uint32(rnd), np.uint32(ka), np.uint32(kb), 0xA511E9B3, 0x5851F42D)
            l, r = r, l ^ (c0 & mask)
        x = (l << np.uint32(half)) | r
        ok = x < n
        out[pending[ok]] = x[ok]
        pending, x = pending[~ok], x[~ok]
    return out.astype(np.int64)


def shuffle(groups, idx_in, idx_out, epoch, active, seed):
    i_in, i_out, ep = _np(idx_in), _np(idx_out), _np(epoch)
    act = _np(active)
    s0, s1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    for gi, g in enumerate(groups):
        if act is not None and not act[gi]:
            continue
        n, off = int(g["n_rows"]), int(g["idx_off"])
        ka = (s0 ^ ((int(g["client"]) * 0x9E3779B1) & 0xFFFFFFFF)) & 0xFFFFFFFF   # by client index: independent of the packing
        kb = (s1 + int(ep[gi])) & 0xFFFFFFFF
        i_out[off: off + n] = i_in[off: off + n][_feistel_perm(n, ka, kb)]


def fwd_bwd(groups, tiles, params, X, ldx, y, idx, batch, scratch, step, active, fed_state):
    """Full-batch gradient of each group; written to scratch record 0 (records 1.. stay 0)."""
    p, Xn, yn, idn, sc, stp = _np(params), _np(X), _np(y), _np(idx), _np(scratch), _np(step)
    act = _np(active)
    st = _np(fed_state).view(FEDSTATE_DTYPE) if fed_state is not None else None
    for gi, g in enumerate(groups):
        if act is not None and not act[gi]:
            continue
        if st is not None and st[int(g["fed"])]["stop"]:
            continue
        bb = _batch_bounds(g, batch)
        if bb is None:
            continue
        b0, nb = bb
        stp[gi] += 1
        L, dims, w_off, b_off, ld, end = _layout(g)
        P = int(g["slab_size"])
        slab = p[int(g["slab_off"]): int(g["slab_off"]) + P]
        rows = idn[int(g["idx_off"]) + b0: int(g["idx_off"]) + b0 + nb]
        Xb, yb = Xn[rows], yn[rows].astype(np.int64)
        a = _forward(g, slab, Xb)
        z = a[-1]
        inv_n = F32(1.0 / nb)
        if int(g["head"]) == 0:
            mx = z.max(axis=1, keepdims=True)
            lse = mx[:, 0] + np.log(np.exp(z - mx).sum(axis=1))
            loss_sum = float((lse - z[np.arange(nb), yb]).sum())
            dz = np.exp(z - lse[:, None]).astype(F32)
            dz[np.arange(nb), yb] -= F32(1)
            dz *= inv_n
        else:
            pz = (F32(1) / (F32(1) + np.exp(-z[:, 0]))).astype(F32)
            eps = F32(1.1920929e-07)
            pc = np.clip(pz, eps, F32(1) - eps)
            yt = yb.astype(F32)
            loss_sum = float(-(yt * np.log(pc) + (F32(1) - yt) * np.log(F32(1) - pc)).sum())
            dz = ((pz - yt) * inv_n)[:, None].astype(F32)
        stride = P + LOSS_SLOTS
        rec = sc[int(g["scratch_off"]): int(g["scratch_off"]) + stride * int(g["max_tiles"])].reshape(-1, stride)
        rec[...] = 0
        for l in range(L - 1, -1, -1):
            x = a[l]
            if x.shape[1] < ld[l]:
                x = np.pad(x, ((0, 0), (0, ld[l] - x.shape[1])))
            rec[0, w_off[l]: w_off[l] + dims[l + 1] * ld[l]] = (dz.T @ x).astype(F32).reshape(-1)
            rec[0, b_off[l]: b_off[l] + dims[l + 1]] = dz.sum(axis=0)
            if l > 0:
                W = slab[w_off[l]: w_off[l] + dims[l + 1] * ld[l]].reshape(dims[l + 1], ld[l])[:, : dims[l]]
                dz = ((dz @ W) * _act_bwd(a[l][:, : dims[l]], int(g["act"]))).astype(F32)
        rec[0, P] = F32(loss_sum)


def step_scalars(g, t):
    lr = float(g["lr"])
    if int(g["lr_step_size"]) > 0:
        lr *= float(g["lr_gamma"]) ** ((t - 1) // int(g["lr_step_size"]))
    opt = int(g["opt"])
    if opt == 1:
        bc1, bc2 = 1.0 - float(g["beta1"]) ** t, 1.0 - float(g["beta2"]) ** t
        return F32(lr / bc1), F32(1.0 / math.sqrt(bc2))
    if opt == 2:
        bc1, bc2 = 1.0 - float(g["beta1"]) ** t, 1.0 - float(g["beta2"]) ** t
        return F32(lr * math.sqrt(bc2) / bc1), F32(1)
    return F32(lr), F32(1)


def batch_stats_stride(max_slab):
    return 2 + (max_slab // 4 + 255) // 256


def reduce_opt(groups, max_slab, params, m, v, scratch, batch, step, active, fed_state, loss_out, batch_stats, max_batches):
    p, mn, vn, sc, stp = _np(params), _np(m), _np(v), _np(scratch), _np(step)
    act, lo, bs_arr = _np(active), _np(loss_out), _np(batch_stats)
    st = _np(fed_state).view(FEDSTATE_DTYPE) if fed_state is not None else None
    stride_bs = batch_stats_stride(max_slab)
    for gi, g in enumerate(groups):
        if act is not None and not act[gi]:
            continue
        if st is not None and st[int(g["fed"])]["stop"]:
            continue
        bb = _batch_bounds(g, batch)
        if bb is None:
            continue
        b0, nb = bb
        L, dims, w_off, b_off, ld, end = _layout(g)
        P, off = int(g["slab_size"]), int(g["slab_off"])
        stride = P + LOSS_SLOTS
        ntiles = (nb + TILE_ROWS - 1) // TILE_ROWS
        rec = sc[int(g["scratch_off"]): int(g["scratch_off"]) + stride * int(g["max_tiles"])].reshape(-1, stride)[:ntiles]
        grad = np.zeros(P, dtype=F32)
        for t in range(ntiles):
            grad += rec[t, :P]
        w = p[off: off + P]
        is_w = np.zeros(P, dtype=bool)
        for l in range(L):
            is_w[w_off[l]: b_off[l]] = True
        sumsq = float((w[is_w].astype(np.float64) ** 2).sum())
        alpha = F32(g["alpha"])
        if alpha != 0:
            grad[is_w] += (alpha / F32(nb)) * w[is_w]
        lr_eff, isb2 = step_scalars(g, int(stp[gi]))
        b1, b2, eps = F32(g["beta1"]), F32(g["beta2"]), F32(g["eps"])
        opt = int(g["opt"])
        sl = slice(off, off + end)
        gr = grad[:end]
        if opt == 1:
            mn[sl] = b1 * mn[sl] + (F32(1) - b1) * gr
            vn[sl] = b2 * vn[sl] + (F32(1) - b2) * gr * gr
            p[sl] = p[sl] - lr_eff * (mn[sl] / (np.sqrt(vn[sl]) * isb2 + eps))
        elif opt == 2:
            mn[sl] = b1 * mn[sl] + (F32(1) - b1) * gr
            vn[sl] = b2 * vn[sl] + (F32(1) - b2) * gr * gr
            p[sl] = p[sl] - lr_eff * mn[sl] / (np.sqrt(vn[sl]) + eps)
        else:
            mom = F32(g["momentum"])
            vel = (mn[sl] if mn is not None else np.zeros(end, dtype=F32))
            vel = mom * vel - lr_eff * gr
            upd = (mom * vel - lr_eff * gr) if int(g["nesterov"]) else vel
            if mn is not None:
                mn[sl] = vel
            p[sl] = p[sl] + upd
        ls = float(sum(float(rec[t, P]) for t in range(ntiles)))
        if lo is not None:
            lo[gi] = F32(ls / nb)
        if bs_arr is not None:
            row = bs_arr.reshape(-1)[(gi * max_batches + batch) * stride_bs: (gi * max_batches + batch + 1) * stride_bs]
            row[...] = 0
            row[0], row[1], row[2] = ls, nb, sumsq


def epoch_end(groups, batch_stats, max_batches, max_slab, tol, n_iter_no_change, max_iter, active, epoch, no_improve, best_loss, last_loss):
    bs_arr, act, ep, ni, bl, ll = (_np(t) for t in (batch_stats, active, epoch, no_improve, best_loss, last_loss))
    stride_bs = batch_stats_stride(max_slab)
    flat = bs_arr.reshape(-1)
    for gi, g in enumerate(groups):
        if not act[gi]:
            continue
        n = int(g["n_rows"])
        bs = min(int(g["batch_size"]), n) if int(g["batch_size"]) > 0 else n
        nbatches = (n + bs - 1) // bs
        acc = 0.0
        for b in range(nbatches):
            row = flat[(gi * max_batches + b) * stride_bs: (gi * max_batches + b + 1) * stride_bs]
            acc += row[0] + 0.5 * float(g["alpha"]) * row[2:].sum()
        loss = acc / n
        ll[gi] = loss
        ep[gi] += 1
        if loss > bl[gi] - tol:
            ni[gi] += 1
        else:
            ni[gi] = 0
        if loss < bl[gi]:
            bl[gi] = loss
        if ni[gi] > n_iter_no_change or ep[gi] >= max_iter:
            act[gi] = 0


def predict(g, slab, Xb):
    z = _forward(g, slab, Xb)[-1]
    if int(g["head"]) == 0:
        return z.argmax(axis=1)
    return (z[:, 0] > 0).astype(np.int64)


def evaluate(groups, jobs, tiles, params, global_, X, ldx, y, idx, conf, fed_state):
    p, gl, Xn, yn, idn, cf = _np(params), _np(global_), _np(X), _np(y), _np(idx), _np(conf)
    st = _np(fed_state).view(FEDSTATE_DTYPE) if fed_state is not None else None
    cf = cf.reshape(-1, MAX_CLASSES, MAX_CLASSES)
    for j in jobs:
        g = groups[int(j["group"])]
        if st is not None and st[int(g["fed"])]["stop"]:
            continue
        src = gl if int(j["from_global"]) else p
        slab = src[int(j["slab_off"]): int(j["slab_off"]) + int(g["slab_size"])]
        rows = idn[int(j["idx_off"]): int(j["idx_off"]) + int(j["n_rows"])]
        pred = predict(g, slab, Xn[rows])
        np.add.at(cf[int(j["conf_slot"])], (yn[rows].astype(np.int64), pred), 1)


def _metrics(cm, C):
    cm = cm[:C, :C].astype(np.float64)
    total = cm.sum()
    if total == 0:
        return np.zeros(4)
    tp, sup, prd = np.diag(cm), cm.sum(1), cm.sum(0)
    with np.errstate(divide="ignore", invalid="ignore"):
        pr = np.where(prd > 0, tp / prd, 0.0)
        rc = np.where(sup > 0, tp / sup, 0.0)
        f1 = np.where(sup + prd > 0, 2 * tp / (sup + prd), 0.0)
    w = sup / total
    return np.array([tp.sum() / total, (pr * w).sum(), (rc * w).sum(), (f1 * w).sum()])


def fedavg(groups, feds, fed_state, max_slab, params, global_, conf, hist_local, hist_global, write_back, do_metrics,
           begin_next=False):
    p, gl, cf = _np(params), _np(global_), _np(conf)
    hl, hg = _np(hist_local), _np(hist_global)
    st = _np(fed_state).view(FEDSTATE_DTYPE)
    n_groups = len(groups)
    for f, F in enumerate(feds):
        if st[f]["stop"]:
            continue
        P = int(F["slab_size"])
        acc = np.zeros(P, dtype=F32)
        members = range(int(F["first_group"]), int(F["first_group"]) + int(F["n_clients"]))
        for gi in members:
            g = groups[gi]
            w = p[int(g["slab_off"]): int(g["slab_off"]) + P]
            acc += (w * F32(g["agg_num"])) / F32(g["agg_den"])
        gl[int(F["global_off"]): int(F["global_off"]) + P] = acc
        if write_back:
            for gi in members:
                g = groups[gi]
                p[int(g["slab_off"]): int(g["slab_off"]) + P] = acc
        if do_metrics:
            C = int(F["n_classes"])
            cms = cf.reshape(-1, MAX_CLASSES, MAX_CLASSES)
            r = int(st[f]["round"]) - 1
            loc = [_metrics(cms[gi], C) for gi in members]
            if hl is not None and 0 <= r < int(F["max_rounds"]):
                for gi, mloc in zip(members, loc):
                    hl.reshape(-1, n_groups, 4)[r, gi] = mloc
            if int(F["metric_mode"]) == 1:
                glob = _metrics(np.sum([cms[gi] for gi in members], axis=0), C)
            else:
                glob = np.sum(loc, axis=0) / len(loc)
            st[f]["last"] = glob
            if hg is not None and 0 <= r < int(F["max_rounds"]):
                hg.reshape(len(feds), -1, 4)[f, r] = glob
            if int(F["patience"]) > 0:
                close = bool(st[f]["has_anchor"]) and bool(np.all(
                    np.abs(glob - st[f]["anchor"]) <= float(F["atol"]) + float(F["rtol"]) * np.abs(st[f]["anchor"])))
                if close:
                    st[f]["count"] -= 1
                    if st[f]["count"] == 0:
                        st[f]["stop_pending"] = 1
                else:
                    st[f]["anchor"] = glob
                    st[f]["has_anchor"] = 1
                    st[f]["count"] = int(F["patience"])
            st[f]["done"] = st[f]["round"]
        if begin_next:
            if st[f]["stop_pending"]:
                st[f]["stop"] = 1
            if not st[f]["stop"]:
                st[f]["round"] += 1
            first, n = int(F["first_group"]), int(F["n_clients"])
            cf.reshape(-1, MAX_CLASSES, MAX_CLASSES)[first: first + n] = 0
