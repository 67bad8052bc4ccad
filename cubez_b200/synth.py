"""Synthetic start fields that are a function of the GLOBAL node index only.

bench.py and the parity tests need a non-trivial field (the reference problem's own start, P = 0,
keeps most cells at exactly zero for the first iterations) that is IDENTICAL for every
decomposition of the same grid, so that a run on N j-slabs can be compared with the single-domain
run, which is what the reference's result is defined as (cz_Poisson.cpp:159-235 with numProc = 1).
The value of node (k, i, j) (Fortran global indices, 1-based) is a 32-bit integer hash of the index
mapped to [-1, 1) on a 2^-23 grid -- exact in float and double, so the torch (device) and numpy
(host, used with the CPU oracle in tests) versions below agree bit for bit.
"""
import numpy as np

_M32 = 0xFFFFFFFF


def _mix(h, xp_and):
    # h: int64 array-like holding values < 2^36; only the low 32 bits of each product are kept
    h = xp_and((h ^ (h >> 17)) * 0x9E3779B1, _M32)
    h = h ^ (h >> 15)
    h = xp_and(h * 0x85EBCA77, _M32)
    h = h ^ (h >> 13)
    return h >> 8            # 24 bits


def planes_numpy(jg0, nj, ni, nk, dtype=np.float32):
    """Values of global planes jg0 .. jg0+nj-1, nodes i = 1..ni, k = 1..nk -> array (nj, ni, nk)."""
    j = np.arange(jg0, jg0 + nj, dtype=np.int64).reshape(-1, 1, 1)
    i = np.arange(1, ni + 1, dtype=np.int64).reshape(1, -1, 1)
    k = np.arange(1, nk + 1, dtype=np.int64).reshape(1, 1, -1)
    h = _mix((j << 24) | (i << 12) | k, lambda a, m: a & m)
    return (h.astype(np.float64) * 2.0 ** -23 - 1.0).astype(dtype)


def planes_torch(jg0, nj, ni, nk, dtype, device):
    import torch
    j = torch.arange(jg0, jg0 + nj, dtype=torch.int64, device=device).view(-1, 1, 1)
    i = torch.arange(1, ni + 1, dtype=torch.int64, device=device).view(1, -1, 1)
    k = torch.arange(1, nk + 1, dtype=torch.int64, device=device).view(1, 1, -1)
    h = _mix((j << 24) | (i << 12) | k, lambda a, m: a & m)
    return (h.to(torch.float64) * 2.0 ** -23 - 1.0).to(dtype)


def fill(field, sz, head_j=1, chunk=32, gny=None):
    """Fill nodes 1..sz of a (sz[1]+4, sz[0]+4, sz[2]+4) field (torch CUDA tensor or numpy array; C index =
    Fortran index + 1) with the hash of the GLOBAL index; head_j = global j of local plane 1 (DomainInfo head).
    gny (global node count along j) given: the two guard planes on each j side that exist in the global grid are
    filled too, i.e. a slab starts with the halo an exchange would deliver.
    Boundary nodes are overwritten too -- call bc_k afterwards, as the reference does after every iteration."""
    assert max(sz) < 4096 and sz[1] + head_j < (1 << 20)
    ni, nj, nk = int(sz[0]), int(sz[1]), int(sz[2])
    is_np = isinstance(field, np.ndarray)
    lo, hi = 1, nj                                    # local Fortran planes to fill
    if gny is not None:
        lo = max(-1, 1 - (head_j - 1))                # global plane 1 is the first that exists
        hi = min(nj + 2, int(gny) - (head_j - 1))
    for j0 in range(lo, hi + 1, chunk):
        m = min(chunk, hi + 1 - j0)
        jg0 = j0 + head_j - 1
        if is_np:
            field[j0 + 1:j0 + 1 + m, 2:2 + ni, 2:2 + nk] = planes_numpy(jg0, m, ni, nk, field.dtype)
        else:
            field[j0 + 1:j0 + 1 + m, 2:2 + ni, 2:2 + nk] = planes_torch(jg0, m, ni, nk, field.dtype, field.device)
    return field


def checksum_owned(field, idx):
    """Order-independent exact checksum of the owned inner cells (Fortran ranges idx = ist,ied,jst,jed,kst,ked):
    the sum of the IEEE bit patterns as signed integers, modulo 2^64.  Sums over slabs add up to the
    single-domain value, so ranks can all-reduce it."""
    ist, ied, jst, jed, kst, ked = [int(v) for v in idx]
    sub = field[jst + 1:jed + 2, ist + 1:ied + 2, kst + 1:ked + 2]
    if isinstance(field, np.ndarray):
        it = np.int32 if field.dtype == np.float32 else np.int64
        return int(np.ascontiguousarray(sub).view(it).astype(np.int64).sum(dtype=np.int64))
    import torch
    it = torch.int32 if field.dtype == torch.float32 else torch.int64
    tot = 0
    for a in range(0, sub.shape[0], 64):      # bounded temporaries
        tot += int(sub[a:a + 64].contiguous().view(it).to(torch.int64).sum().item())
    tot &= (1 << 64) - 1
    return tot - (1 << 64) if tot >= (1 << 63) else tot
