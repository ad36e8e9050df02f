"""Seeded synthetic inputs of the shapes BASELINE.json names: corner-rich grayscale frames,
stereo partners with a piecewise-constant disparity field, and the map/frame structures the matchers consume.
Pure numpy; identical bytes for identical seeds on every machine (PCG64)."""
import numpy as np


def _value_noise(rng, h, w, cell, amp):
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.random((gh, gw), dtype=np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = ys.astype(np.int32)
    x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)


def image(seed, w, h, n_rect=90, n_tri=50, noise_sigma=2.0):
    """I(seed, w, h): 4 octaves of bilinear value noise + filled rectangles and triangles + Gaussian noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    img = np.full((h, w), 38.0, np.float32)
    for cell, amp in ((64, 96.0), (32, 48.0), (16, 24.0), (8, 12.0)):
        img += _value_noise(rng, h, w, cell, amp)
    for _ in range(n_rect):
        x0, y0 = int(rng.integers(0, w - 8)), int(rng.integers(0, h - 8))
        rw, rh = int(rng.integers(6, w // 6)), int(rng.integers(6, h // 6))
        img[y0:y0 + rh, x0:x0 + rw] = float(rng.integers(0, 256))
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(n_tri):
        cx, cy = rng.integers(0, w), rng.integers(0, h)
        p = np.stack([cx + rng.integers(-w // 10, w // 10, 3), cy + rng.integers(-h // 8, h // 8, 3)], 1).astype(np.float32)
        x0, x1 = int(max(p[:, 0].min(), 0)), int(min(p[:, 0].max() + 1, w))
        y0, y1 = int(max(p[:, 1].min(), 0)), int(min(p[:, 1].max() + 1, h))
        if x1 <= x0 or y1 <= y0:
            continue
        X = xx[y0:y1, x0:x1].astype(np.float32)
        Y = yy[y0:y1, x0:x1].astype(np.float32)

        def edge(a, b):
            return (b[0] - a[0]) * (Y - a[1]) - (b[1] - a[1]) * (X - a[0])

        e0, e1, e2 = edge(p[0], p[1]), edge(p[1], p[2]), edge(p[2], p[0])
        inside = ((e0 >= 0) & (e1 >= 0) & (e2 >= 0)) | ((e0 <= 0) & (e1 <= 0) & (e2 <= 0))
        img[y0:y1, x0:x1][inside] = float(rng.integers(0, 256))
    img += rng.normal(0.0, noise_sigma, (h, w)).astype(np.float32)
    return np.clip(img + 0.5, 0, 255).astype(np.uint8)


def stereo_right(left, seed, d_min=4.0, d_max=40.0, noise_sigma=1.0):
    """Right view: the left image shifted by a piecewise-constant disparity (3 depth planes laid out as vertical
    bands with seeded borders), bilinear in x, plus independent sensor noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    h, w = left.shape
    cuts = np.sort(rng.integers(w // 6, w - w // 6, 2))
    disp = np.empty(w, np.float32)
    d = rng.uniform(d_min, d_max, 3).astype(np.float32)
    disp[:cuts[0]] = d[0]
    disp[cuts[0]:cuts[1]] = d[1]
    disp[cuts[1]:] = d[2]
    xs = np.arange(w, dtype=np.float32) + disp  # right(x) = left(x + d)
    x0 = np.clip(np.floor(xs).astype(np.int32), 0, w - 1)
    x1 = np.clip(x0 + 1, 0, w - 1)
    fx = (xs - np.floor(xs))[None, :]
    L = left.astype(np.float32)
    r = L[:, x0] * (1 - fx) + L[:, x1] * fx
    r += rng.normal(0.0, noise_sigma, (h, w)).astype(np.float32)
    return np.ascontiguousarray(np.clip(r + 0.5, 0, 255).astype(np.uint8))


def adversarial(kind, w, h, seed=0):
    """Parity-only inputs: constant (no keypoints), 8-px checkerboard (response ties), uniform noise."""
    if kind == "constant":
        return np.full((h, w), 117, np.uint8)
    if kind == "checker":
        yy, xx = np.mgrid[0:h, 0:w]
        return (((yy // 8 + xx // 8) % 2) * 255).astype(np.uint8)
    if kind == "noise":
        return np.random.Generator(np.random.PCG64(seed)).integers(0, 256, (h, w), dtype=np.uint8)
    raise ValueError(kind)


SHAPES = {"euroc": (752, 480), "kitti": (1241, 376), "tum": (640, 480)}


# ---------------------------------------------------------------------------------------------------------------
# Matcher inputs. Everything below is plain numpy on host arrays.
# ---------------------------------------------------------------------------------------------------------------
KP64 = np.dtype([("x", "<f8"), ("y", "<f8"), ("size", "<f8"), ("angle", "<f8"), ("response", "<f8"),
                 ("octave", "<i4"), ("_pad", "<i4")])


def feature_grid(kps64, w, h, cell=20):
    """Saiga::FeatureGridBounds2<double,20>::computeFromIntrinsicsDist for a distortion-free camera (bounding box of
    the image border +-0.1) and FeatureGrid2::create (counting sort into row-major cells; the permutation keeps the
    input order inside a cell). Returns (bounds dict, cells int32 [rows*cols,2], permutation)."""
    bmin = np.array([0.0, 0.0]) - 0.1
    bmax = np.array([w - 1.0, h - 1.0]) + 0.1
    inv = 1.0 / cell
    rows = int(np.ceil((bmax[1] - bmin[1]) / cell))
    cols = int(np.ceil((bmax[0] - bmin[0]) / cell))
    px = (kps64["x"] - bmin[0]) * inv
    py = (kps64["y"] - bmin[1]) * inv
    cx = np.floor(px).astype(np.int64)
    cy = np.floor(py).astype(np.int64)
    inside = (kps64["x"] >= bmin[0]) & (kps64["y"] >= bmin[1]) & (kps64["x"] < bmax[0]) & (kps64["y"] < bmax[1])
    cid = np.where(inside, cy * cols + cx, rows * cols)
    order = np.argsort(cid, kind="stable")
    perm = np.empty(len(kps64), np.int32)
    perm[order] = np.arange(len(kps64), dtype=np.int32)
    counts = np.bincount(cid[inside], minlength=rows * cols)[: rows * cols]
    begin = np.concatenate([[0], np.cumsum(counts)[:-1]])
    cells = np.stack([begin, begin + counts], 1).astype(np.int32)
    bounds = dict(bmin_x=bmin[0], bmin_y=bmin[1], bmax_x=bmax[0], bmax_y=bmax[1], inv=inv, rows=rows, cols=cols)
    return bounds, cells, perm


def apply_permutation(perm, *arrays):
    out = []
    for a in arrays:
        b = np.empty_like(a)
        b[perm] = a
        out.append(b)
    return out


def flip_bits(desc, rng, max_flips):
    """Copies of 256-bit descriptors with 0..max_flips random bits flipped."""
    d = desc.copy().view(np.uint8).reshape(len(desc), 32)
    for i in range(len(d)):
        for b in rng.integers(0, 256, int(rng.integers(0, max_flips + 1))):
            d[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return d.view(np.uint64).reshape(len(desc), 4)


def quat_from_axis_angle(axis, angle):
    axis = np.asarray(axis, np.float64)
    axis = axis / np.linalg.norm(axis)
    s = np.sin(angle / 2)
    return np.array([axis[0] * s, axis[1] * s, axis[2] * s, np.cos(angle / 2)])


def map_points_from_keypoints(kps64, desc, K, rng, n_points, scale_table, max_flips=20, depth_range=(3.0, 60.0)):
    """Local-map points back-projected from (a resampling of) the frame's own keypoints at seeded depths, the
    descriptor a noisy copy of the keypoint's. Returns dict of arrays (pos, normal, desc, ref_depth, ref_level,
    octave, angle, src)."""
    fx, fy, cx, cy = K
    src = rng.integers(0, len(kps64), n_points)
    jitter = rng.normal(0, 0.7, (n_points, 2))
    u = kps64["x"][src] + jitter[:, 0]
    v = kps64["y"][src] + jitter[:, 1]
    depth = rng.uniform(depth_range[0], depth_range[1], n_points)
    pos = np.stack([(u - cx) / fx * depth, (v - cy) / fy * depth, depth], 1)
    normal = -pos / np.linalg.norm(pos, axis=1, keepdims=True)  # facing the camera at the origin
    lvl = kps64["octave"][src].astype(np.int32)
    # consistent with the scale prediction: seen at `depth` on level `lvl`
    return dict(pos=pos, normal=normal, desc=flip_bits(desc[src], rng, max_flips), ref_depth=depth.astype(np.float32),
                ref_level=lvl, octave=lvl, angle=(kps64["angle"][src] + rng.normal(0, 4, n_points)).astype(np.float32),
                src=src.astype(np.int32))


def bow_feature_vector(desc, rng, n_nodes=100, bits=7):
    """A FeatureVector-shaped grouping (MiniBow2.h:98-132) without the vocabulary blob: the node of a descriptor is
    a hash of its leading bits (similar descriptors share nodes), node ids ascending, feature order inside a node
    shuffled like the std::sort in FeatureVector::setFeatures leaves it. Returns (node_ids, offsets, indices)."""
    key = (desc[:, 0] & np.uint64((1 << bits) - 1)).astype(np.int64) % n_nodes
    ids = np.unique(key)
    offsets = [0]
    indices = []
    for nid in ids:
        members = np.nonzero(key == nid)[0]
        rng.shuffle(members)
        indices.append(members)
        offsets.append(offsets[-1] + len(members))
    return ids.astype(np.int32) * 3 + 11, np.array(offsets, np.int32), np.concatenate(indices).astype(np.int32)


def vocabulary_arrays(seed, k, L, n_train=4000, zero_weight_fraction=0.1, ragged=True):
    """A random but well-formed vocabulary tree in loadRaw's node order (depth-first creation as HKmeansStep numbers
    the nodes): each node's children descriptors are the parent's with bits flipped, so a descent is meaningful.
    ragged: some inner nodes get fewer than k children and some branches end early (create() produces both).
    Returns (k, L, parent, weight, word_id, desc)."""
    rng = np.random.default_rng(seed)
    parent, weight, word_id, desc, level = [0], [0.0], [0], [rng.integers(0, 2**63, 4).astype(np.uint64)], [0]

    def grow(node, lvl):
        # HKmeansStep (MiniBow2.h:508-700): the k children are appended first, then each is expanded in turn
        nk = k if not ragged or rng.random() > 0.15 else int(rng.integers(1, k + 1))
        kids = []
        for _ in range(nk):
            d = desc[node].copy()
            for _ in range(max(2, 96 >> lvl)):
                d[rng.integers(0, 4)] ^= np.uint64(1) << np.uint64(rng.integers(0, 64))
            parent.append(node), weight.append(0.0), word_id.append(0), desc.append(d), level.append(lvl + 1)
            kids.append(len(parent) - 1)
        if lvl + 1 < L:
            for c in kids:
                if not ragged or rng.random() > 0.05:
                    grow(c, lvl + 1)

    grow(0, 0)
    n = len(parent)
    has_child = np.zeros(n, bool)
    has_child[np.array(parent[1:])] = True
    nw = 0
    for i in range(1, n):  # createWords (:799-819): leaves numbered in node order
        if not has_child[i]:
            word_id[i] = nw
            nw += 1
            weight[i] = 0.0 if rng.random() < zero_weight_fraction else float(np.log(rng.uniform(1.05, 40.0)))
    return k, L, np.array(parent, np.int32), np.array(weight, np.float64), np.array(word_id, np.int32), np.stack(desc)


def vocabulary_arrays_bfs(seed, k, L, zero_weight_fraction=0.02):
    """A full k-ary vocabulary of depth L (ORBvoc: k=10, L=6, 1 111 111 nodes) numbered breadth-first, built with
    array operations only: every child's descriptor is its parent's with a sparse random mask XORed in, sparser on
    deeper levels. Returns (k, L, parent, weight, word_id, desc)."""
    rng = np.random.default_rng(seed)
    parent, desc = [np.zeros(1, np.int64)], [rng.integers(0, 2**63, (1, 4)).astype(np.uint64)]
    start = 0
    for lvl in range(1, L + 1):
        prev = desc[-1]
        n = len(prev) * k
        par = start + np.arange(n) // k
        mask = rng.integers(0, 2**63, (n, 4)).astype(np.uint64) << np.uint64(1) | rng.integers(0, 2, (n, 4)).astype(np.uint64)
        for _ in range(min(lvl, 4)):  # bit density 1/2, 1/4, 1/8, ... with depth
            mask &= rng.integers(0, 2**63, (n, 4)).astype(np.uint64) << np.uint64(1) | rng.integers(0, 2, (n, 4)).astype(np.uint64)
        desc.append(prev[np.arange(n) // k] ^ mask)
        parent.append(par)
        start += len(prev)
    parent = np.concatenate(parent).astype(np.int32)
    desc = np.concatenate(desc)
    n = len(parent)
    n_leaf = k**L
    weight = np.zeros(n, np.float64)
    word_id = np.zeros(n, np.int32)
    w = np.log(rng.uniform(1.05, 40.0, n_leaf))
    w[rng.random(n_leaf) < zero_weight_fraction] = 0.0
    weight[n - n_leaf:] = w
    word_id[n - n_leaf:] = np.arange(n_leaf)
    return k, L, parent, weight, word_id, desc
