"""Pure-Python mirror of czb_slab_decompose / czb_brick_decompose (cubez_b200/csrc/comm.cpp): 1-D slab decomposition
of the cube along j, the slowest axis, and general 3-D bricks (§8 f-4).  Kept bit-for-bit in sync with the C++
version by tests/test_decomp.py."""

GUIDE = 2  # reference src/cz_cpp/cz_Define.h:40


def slab_decompose(gsz, nranks, rank):
    """Return dict(sz, head, nID, idx) for `rank` of `nranks`; raises ValueError if the cube
    has fewer inner j-planes than ranks."""
    n = gsz[1] - 2
    if nranks < 1 or rank < 0 or rank >= nranks or n < nranks:
        raise ValueError("cannot cut %d inner planes into %d slabs" % (n, nranks))
    base, rem = divmod(n, nranks)
    cnt = base + (1 if rank < rem else 0)
    first = 2 + rank * base + min(rank, rem)
    lo_edge, hi_edge = rank == 0, rank == nranks - 1
    sz = [gsz[0], cnt + int(lo_edge) + int(hi_edge), gsz[2]]
    head = [1, 1 if lo_edge else first, 1]
    nID = [-1, -1, -1 if lo_edge else rank - 1, -1 if hi_edge else rank + 1, -1, -1]
    idx = [2, sz[0] - 1, 2 if lo_edge else 1, sz[1] - 1 if hi_edge else sz[1], 2, sz[2] - 1]
    return dict(sz=sz, head=head, nID=nID, idx=idx)


def brick_decompose(gsz, gdv, rank):
    """Mirror of czb_brick_decompose: gdv = divisions along x (i), y (j), z (k); rank = ri + gdv[0]*(rj + gdv[1]*rk)."""
    nranks = gdv[0] * gdv[1] * gdv[2]
    if min(gdv) < 1 or rank < 0 or rank >= nranks:
        raise ValueError("bad division %r / rank %d" % (gdv, rank))
    r3 = [rank % gdv[0], (rank // gdv[0]) % gdv[1], rank // (gdv[0] * gdv[1])]
    stride = [1, gdv[0], gdv[0] * gdv[1]]
    sz, head, nID, idx = [0] * 3, [0] * 3, [-1] * 6, [0] * 6
    for a in range(3):
        n, parts, r = gsz[a] - 2, gdv[a], r3[a]
        if n < parts:
            raise ValueError("cannot cut %d inner nodes into %d parts" % (n, parts))
        base, rem = divmod(n, parts)
        cnt = base + (1 if r < rem else 0)
        first = 2 + r * base + min(r, rem)
        lo_edge, hi_edge = r == 0, r == parts - 1
        sz[a] = cnt + int(lo_edge) + int(hi_edge)
        head[a] = 1 if lo_edge else first
        nID[2 * a] = -1 if lo_edge else rank - stride[a]
        nID[2 * a + 1] = -1 if hi_edge else rank + stride[a]
        idx[2 * a] = 2 if lo_edge else 1
        idx[2 * a + 1] = sz[a] - 1 if hi_edge else sz[a]
    return dict(sz=sz, head=head, nID=nID, idx=idx)


def inner_cells(idx):
    return (idx[1] - idx[0] + 1) * (idx[3] - idx[2] + 1) * (idx[5] - idx[4] + 1)


def field_shape(sz, g=GUIDE):
    """numpy/torch shape of a field: C order (j, i, k), k contiguous = Fortran (k, i, j)."""
    return (sz[1] + 2 * g, sz[0] + 2 * g, sz[2] + 2 * g)
