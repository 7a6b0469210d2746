"""ctypes access to oracle/liboracle.so (the C restatement).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import torch

_HERE = Path(__file__).resolve().parent
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = _HERE / "liboracle.so"
        if not so.exists() or so.stat().st_mtime < (_HERE / "nkbs_oracle.c").stat().st_mtime:
            subprocess.run(["make", "-s", "-C", str(_HERE)], check=True)
        _lib = C.CDLL(str(so))
    return _lib


def _f(t):
    return None if t is None else t.detach().contiguous().float().numpy()


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def positions(mode, contraction, aabb, x=None, origins=None, directions=None, starts=None, ends=None):
    if mode == 2:
        R, S = starts.shape
    else:
        R, S = x.shape[0], 1
    xn = np.empty((R * S, 3), np.float32)
    sel = np.empty((R * S,), np.float32)
    a = [_f(t) for t in (x, origins, directions, starts, ends)]
    ab = np.asarray(aabb, np.float32)
    lib().oracle_positions(C.c_int(mode), *[_p(v) for v in a], C.c_int64(R), C.c_int32(S), C.c_int(int(contraction)),
                           _p(ab), _p(xn), _p(sel))
    return torch.from_numpy(xn), torch.from_numpy(sel)


def hash_encode(xn, table, scalings, log2_T, want_features=True):
    xn_, tb = _f(xn), _f(table)
    sc = np.asarray(scalings, np.float32)
    N, L = xn_.shape[0], len(sc)
    feats = np.empty((N, 2 * L), np.float32) if want_features else None
    idx = np.empty((N, L, 8), np.int64)
    lib().oracle_hash_encode(_p(xn_), C.c_int64(N), _p(tb), _p(sc), C.c_int32(L), C.c_int32(log2_T), _p(feats), _p(idx))
    return (torch.from_numpy(feats) if want_features else None), torch.from_numpy(idx)


def initial_sampler(nears, fars, S, t_rand=None):
    n, f = _f(nears.reshape(-1)), _f(fars.reshape(-1))
    R = n.shape[0]
    lin = torch.linspace(0.0, 1.0, S + 1).numpy()
    tr = _f(t_rand)
    per_bin = int(tr is not None and tr.size != R)
    sb = np.empty((R, S + 1), np.float32)
    eb = np.empty((R, S + 1), np.float32)
    lib().oracle_initial_sampler(C.c_int64(R), C.c_int32(S), _p(lin), _p(tr), C.c_int(per_bin), _p(n), _p(f), _p(sb), _p(eb))
    return torch.from_numpy(sb), torch.from_numpy(eb)


def pdf_sample(weights, sbins, nears, fars, num_samples, anneal=1.0, rand=None, hist_pad=0.01, eps=1e-5):
    w, sb = _f(weights), _f(sbins)
    R, S = w.shape
    nb = num_samples + 1
    n, f = _f(nears.reshape(-1)), _f(fars.reshape(-1))
    u = torch.linspace(0.0, 1.0 - (1.0 / nb), steps=nb).numpy()
    rd = _f(rand)
    per_bin = int(rd is not None and rd.size != R)
    ob = np.empty((R, nb), np.float32)
    oe = np.empty((R, nb), np.float32)
    oi = np.empty((R, nb), np.int64)
    lib().oracle_pdf_sample(C.c_int64(R), C.c_int32(S), C.c_int32(nb), _p(w), C.c_float(anneal), C.c_float(hist_pad),
                            C.c_float(eps), _p(sb), _p(u), _p(rd), C.c_int(per_bin), _p(n), _p(f), _p(ob), _p(oe), _p(oi))
    return torch.from_numpy(ob), torch.from_numpy(oe), torch.from_numpy(oi)
