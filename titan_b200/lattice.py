"""Direct-to-array lattice builder (no per-object host handles).

Emits masses and springs in exactly the order of the reference's `Lattice::Lattice`
(src/object.cu:235-296): mass index k + j*nz + i*ny*nz with i outermost; per cell (i, j, k) the seven
(l, m, n) != 0 neighbours in l, m, n order, then the yz / xz anti-diagonals, the three remaining body
diagonals, and finally the xy anti-diagonal; rest length = initial distance (:293-295).  Used by bench.py and
the multi-GPU partitioner, where 27 M `Mass` objects on the host would cost ~40 GB (SURVEY.md hard parts).
tests/test_lattice_builder.py checks it bit for bit against the oracle and the reference's own builder.
"""
import numpy as np

# (left offset, right offset) in (i, j, k); a spring exists when both endpoints are inside the grid
_SLOTS = [((0, 0, 0), (0, 0, 1)), ((0, 0, 0), (0, 1, 0)), ((0, 0, 0), (0, 1, 1)), ((0, 0, 0), (1, 0, 0)),
          ((0, 0, 0), (1, 0, 1)), ((0, 0, 0), (1, 1, 0)), ((0, 0, 0), (1, 1, 1)),
          ((0, 0, 1), (0, 1, 0)), ((0, 0, 1), (1, 0, 0)), ((0, 0, 1), (1, 1, 0)),
          ((1, 0, 1), (0, 1, 0)), ((0, 1, 1), (1, 0, 0)), ((0, 1, 0), (1, 0, 0))]


def lattice_counts(nx, ny, nz):
    a, b, c = int(nx), int(ny), int(nz)
    nm = a * b * c
    ns = ((a - 1) * b * c + a * (b - 1) * c + a * b * (c - 1)
          + 2 * ((a - 1) * (b - 1) * c + (a - 1) * b * (c - 1) + a * (b - 1) * (c - 1))
          + 4 * (a - 1) * (b - 1) * (c - 1))
    return nm, ns


def lattice_positions(center, dims, nx, ny, nz, i_range=None):
    """Mass coordinates in reference order; i_range=(a, b) restricts to the i-planes [a, b) of the nx-plane lattice
    (same doubles as the full lattice: the formula only sees the global index and nx)."""
    def axis(n, d, c, lo=0, hi=None):
        hi = n if hi is None else hi
        u = (np.arange(lo, hi, dtype=np.float64) / (n - 1.0) - 0.5) if n > 1 else np.zeros(1)
        return u * d + c
    lo, hi = (0, nx) if i_range is None else i_range
    x = axis(nx, float(dims[0]), float(center[0]), lo, hi)
    y = axis(ny, float(dims[1]), float(center[1]))
    z = axis(nz, float(dims[2]), float(center[2]))
    nx = hi - lo
    pos = np.empty((nx, ny, nz, 3))
    pos[..., 0] = x[:, None, None]
    pos[..., 1] = y[None, :, None]
    pos[..., 2] = z[None, None, :]
    return pos.reshape(-1, 3)


def lattice_springs(nx, ny, nz, i_planes_per_chunk=8):
    """left[ns], right[ns] (int32) in reference emission order."""
    nm, ns = lattice_counts(nx, ny, nz)
    left = np.empty(ns, dtype=np.int32)
    right = np.empty(ns, dtype=np.int32)
    out = 0
    dims = np.array([nx, ny, nz])
    stride = np.array([ny * nz, nz, 1], dtype=np.int64)
    for i0 in range(0, nx, i_planes_per_chunk):
        i1 = min(nx, i0 + i_planes_per_chunk)
        ii, jj, kk = np.meshgrid(np.arange(i0, i1), np.arange(ny), np.arange(nz), indexing="ij")
        cell = np.stack([ii.ravel(), jj.ravel(), kk.ravel()], axis=1).astype(np.int64)
        ncell = cell.shape[0]
        L = np.empty((ncell, 13), dtype=np.int32)
        R = np.empty((ncell, 13), dtype=np.int32)
        ok = np.empty((ncell, 13), dtype=bool)
        for s, (lo, ro) in enumerate(_SLOTS):
            a = cell + np.array(lo)
            b = cell + np.array(ro)
            ok[:, s] = ((a < dims).all(axis=1)) & ((b < dims).all(axis=1))
            L[:, s] = (a @ stride).astype(np.int32)
            R[:, s] = (b @ stride).astype(np.int32)
        n = int(ok.sum())
        left[out:out + n] = L[ok]
        right[out:out + n] = R[ok]
        out += n
    assert out == ns
    return left, right


def rest_lengths(pos, left, right):
    """Vec::norm of (right.pos - left.pos): sqrt((dx*dx + dy*dy) + dz*dz), include/Titan/vec.h:149-151."""
    d = pos[right] - pos[left]
    return np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2])


def build_lattice(center, dims, nx, ny, nz):
    pos = lattice_positions(center, dims, nx, ny, nz)
    left, right = lattice_springs(nx, ny, nz)
    if left.size <= 50_000_000:
        rest = rest_lengths(pos, left, right)
    else:
        rest = np.empty(left.size)
        for a in range(0, left.size, 20_000_000):
            rest[a:a + 20_000_000] = rest_lengths(pos, left[a:a + 20_000_000], right[a:a + 20_000_000])
    return dict(pos=pos, left=left, right=right, rest=rest)


def brick_dims(nx, ny, nz, tile=128):
    """Brick edge lengths (bx, by, bz), bx*by*bz <= tile, for tiling an nx x ny x nz lattice, minimising the
    modelled bytes per step: every tile streams `tile` mass slots (padding included), one record per spring
    touching it (cut springs are stored by both bricks) and its halo positions."""
    best = None
    for bx in range(1, min(nx, tile) + 1):
        for by in range(1, min(ny, tile // bx) + 1):
            bz = min(nz, tile // (bx * by))
            if bz < 1:
                continue
            tiles = -(-nx // bx) * -(-ny // by) * -(-nz // bz)
            a, b, c = bx - 1, by - 1, bz - 1
            inner = (a * by * bz + bx * b * bz + bx * by * c) + 2 * (a * b * bz + a * by * c + bx * b * c) + 4 * a * b * c
            recs = 26 * bx * by * bz - inner
            halo = (bx + 2) * (by + 2) * (bz + 2) - bx * by * bz
            cost = tiles * (tile * 104 + 16 * recs + 24 * halo)
            key = (cost, bx, by, -bz)
            if best is None or key < best[0]:
                best = (key, (bx, by, bz))
    return best[1]


def _brick_cost(bx, by, bz, tile):
    """Modelled bytes one tile of a bx x by x bz brick streams per step (see brick_dims)."""
    a, b, c = bx - 1, by - 1, bz - 1
    inner = (a * by * bz + bx * b * bz + bx * by * c) + 2 * (a * b * bz + a * by * c + bx * b * c) + 4 * a * b * c
    recs = 26 * bx * by * bz - inner
    halo = (bx + 2) * (by + 2) * (bz + 2) - bx * by * bz
    return tile * 104 + 16 * recs + 24 * halo


def brick_layers(nx, ny, nz, tile=128):
    """Layers of bricks along i: [(thickness, by, bz), ...] with sum(thickness) == nx.

    One brick shape for the whole lattice (brick_dims) wastes the last layer when nx is not a multiple of the brick
    thickness -- e.g. the 37/38-plane slabs of a 300^3 lattice over 8 GPUs need 8 layers of 5 x 5 x 5 bricks (28,800 tiles)
    where 3 layers of 6 x 5 x 4 plus 4 of 5 x 5 x 5 do it in 27,900 (27,000 for 37 planes).  A tile costs about the same
    time whatever its fill, so the layers are chosen by dynamic programming over the thickness, each layer with its
    cheapest cross-section under the byte model of brick_dims.  Same-shaped layers are kept adjacent (thickest first) so
    that congruent bricks share their incidence tables."""
    layer = {}
    for t in range(1, min(nx, tile) + 1):
        best = None
        for by in range(1, min(ny, tile // t) + 1):
            bz = min(nz, tile // (t * by))
            if bz < 1:
                continue
            cost = -(-ny // by) * -(-nz // bz) * _brick_cost(t, by, bz, tile)
            key = (cost, by, -bz)
            if best is None or key < best[0]:
                best = (key, (t, by, bz))
        if best is not None:
            layer[t] = (best[0][0], best[1])
    INF = float("inf")
    cost = [0.0] + [INF] * nx
    pick = [0] * (nx + 1)
    for n in range(1, nx + 1):
        for t, (c, _) in layer.items():
            if t <= n and cost[n - t] + c < cost[n]:
                cost[n], pick[n] = cost[n - t] + c, t
    out, n = [], nx
    while n > 0:
        out.append(layer[pick[n]][1])
        n -= pick[n]
    out.sort(key=lambda d: -d[0])
    uni = brick_dims(nx, ny, nz, tile)
    uniform_cost = -(-nx // uni[0]) * -(-ny // uni[1]) * -(-nz // uni[2]) * _brick_cost(uni[0], uni[1], uni[2], tile)
    if uniform_cost <= cost[nx]:          # nothing to gain: keep the single shape (and its tile numbering)
        return [uni] * (-(-nx // uni[0])) if nx % uni[0] == 0 else None
    return out


def _brick_ids(nx, ny, nz, dims, block):
    bx, by, bz = dims
    ti = np.arange(nx, dtype=np.int64) // bx
    tj = np.arange(ny, dtype=np.int64) // by
    tk = np.arange(nz, dtype=np.int64) // bz
    nti, ntj, ntk = -(-nx // bx), -(-ny // by), -(-nz // bz)
    if block and block > 1 and max(nti, ntj, ntk) > block:
        bi, bj, bk = np.meshgrid(np.arange(nti), np.arange(ntj), np.arange(ntk), indexing="ij")
        keys = (bk % block, bj % block, bi % block, bk // block, bj // block, bi // block)   # last key = most significant
        order = np.lexsort(tuple(k.ravel() for k in keys))
        rank_of = np.empty(order.size, dtype=np.int64)
        rank_of[order] = np.arange(order.size)
        brick_id = rank_of.reshape(nti, ntj, ntk)
    else:
        brick_id = (np.arange(nti)[:, None, None] * ntj + np.arange(ntj)[None, :, None]) * ntk + np.arange(ntk)[None, None, :]
    return brick_id[ti[:, None, None], tj[None, :, None], tk[None, None, :]]


def brick_hint(nx, ny, nz, tile=128, base=0, block=None):
    """Tile id per mass (reference mass order, i outermost) for tb_set_tile_hint: bricks of lattice cells.

    Ids order the tiles in the step kernel.  Bricks are numbered super-block by super-block (`block`^3 bricks each,
    default TB_TILE_BLOCK or 8) rather than plane by plane: the persistent CTAs then work on a compact 3-D
    neighbourhood, so the state a tile gathers as its halo was touched a few hundred tiles ago and is still in L2,
    instead of a whole lattice cross-section (tens of MB of streamed spring records) ago.  When nx is not a multiple of
    the brick thickness the lattice is cut into layers of differently shaped bricks (brick_layers; TB_BRICK_LAYERS=0
    keeps one shape)."""
    import os
    if block is None:
        block = int(os.environ.get("TB_TILE_BLOCK", "8"))
    layers = brick_layers(nx, ny, nz, tile) if os.environ.get("TB_BRICK_LAYERS", "1") != "0" else None
    spec = os.environ.get("TB_BRICK_SPEC")      # experiment knob: "8x4x4*10,5x5x5*4" = layers along i, in this order
    if spec:
        layers = []
        for part in spec.split(","):
            dims, _, rep = part.partition("*")
            layers += [tuple(int(v) for v in dims.split("x"))] * int(rep or 1)
        assert sum(l[0] for l in layers) == nx, "TB_BRICK_SPEC does not add up to nx"
    if not layers or len(set(layers)) == 1:
        return (_brick_ids(nx, ny, nz, brick_dims(nx, ny, nz, tile), block).reshape(-1) + base).astype(np.int32)
    ids = np.empty((nx, ny, nz), dtype=np.int64)
    i0, nxt = 0, 0
    j = 0
    while j < len(layers):                  # one run of equally shaped layers at a time
        run = 1
        while j + run < len(layers) and layers[j + run] == layers[j]:
            run += 1
        planes = layers[j][0] * run
        sub = _brick_ids(planes, ny, nz, layers[j], block)
        ids[i0:i0 + planes] = sub + nxt
        nxt += int(sub.max()) + 1
        i0 += planes
        j += run
    return (ids.reshape(-1) + base).astype(np.int32)
