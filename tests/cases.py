"""Seeded input generators shared by tests/golden/make_golden.py, the parity tests, smoke() and
bench.py.  numpy's legacy RandomState is used because its streams are frozen across numpy
versions, so the committed golden outputs stay valid for regenerated inputs (each fixture also
stores an input checksum that the tests verify)."""
import numpy as np


def level_lengths(t0, n_levels=4):
    """Temporal pyramid of the reference's base encoder: Conv1d(k3,s2,p1) per extra level,
    T_{l+1} = floor((T_l - 1) / 2) + 1   (pdvc/base_encoder.py:37-41)."""
    ts = [int(t0)]
    for _ in range(n_levels - 1):
        ts.append((ts[-1] - 1) // 2 + 1)
    return ts


def level_start_index(shapes2):
    hw = np.asarray(shapes2, dtype=np.int64)
    sizes = hw[:, 0] * hw[:, 1]
    return np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)


def as_shapes2(shapes):
    s = np.asarray(shapes, dtype=np.int64)
    if s.ndim == 1:
        s = np.stack([np.ones_like(s), s], -1)
    return s


def _softmax_lp(z):
    sh = z.shape
    z = z.reshape(sh[0], sh[1], sh[2], -1)
    z = z - z.max(-1, keepdims=True)
    e = np.exp(z)
    return (e / e.sum(-1, keepdims=True)).reshape(sh)


def keep_off_knots(x, sizes_lvl, eps=2e-4):
    """Move pixel coordinates at least `eps` away from integer knots, where grad_loc is
    discontinuous (tolerances): an fp32 floor flip there changes one element by O(1).
    x (N,Lq,M,L,P) normalised; sizes_lvl (L,)."""
    size = np.asarray(sizes_lvl, dtype=np.float64)[None, None, None, :, None]
    pix = x * size - 0.5
    frac = pix - np.floor(pix)
    pix = np.where(frac < eps, np.floor(pix) + eps, pix)
    pix = np.where(frac > 1 - eps, np.floor(pix) + 1 - eps, pix)
    return (pix + 0.5) / size


def make_core_inputs(seed, N, M, D, shapes, Lq, P, lo=-0.05, hi=1.05, two_d=False, off_knots=True):
    """value ~ N(0,1), loc_x ~ U(lo,hi), loc_y = 0.5 (1-D) or ~ U(lo,hi) (2-D),
    attn = softmax(N(0,1)) over L*P, grad_out ~ N(0,1)  (SURVEY 8d, config 1)."""
    rs = np.random.RandomState(seed)
    shapes2 = as_shapes2(shapes)
    L = shapes2.shape[0]
    S = int((shapes2[:, 0] * shapes2[:, 1]).sum())
    value = rs.standard_normal((N, S, M, D))
    x = rs.uniform(lo, hi, (N, Lq, M, L, P))
    if off_knots:
        x = keep_off_knots(x, shapes2[:, 1])
    if two_d:
        y = rs.uniform(lo, hi, (N, Lq, M, L, P))
        if off_knots:
            y = keep_off_knots(y, shapes2[:, 0])
    else:
        y = np.full_like(x, 0.5)
    loc = np.stack([x, y], -1)
    attn = _softmax_lp(rs.standard_normal((N, Lq, M, L, P)))
    grad_out = rs.standard_normal((N, Lq, M * D))
    return dict(value=value, shapes=shapes2, lsi=level_start_index(shapes2), loc=loc, attn=attn,
                grad_out=grad_out)


def make_edge_inputs(seed=7):
    """Edge locations (SURVEY 7 'Tests the build must write'): x exactly 0, 1, 0.5/T, 1-0.5/T,
    integer knots +- 1e-3, far outside, all samples of a query on one row, a level with T = 1,
    one-hot attention rows."""
    rs = np.random.RandomState(seed)
    N, M, D, P = 2, 2, 8, 4
    T = [8, 4, 1]  # powers of two: the exact-knot rows below are exact in fp32 and fp64
    L = len(T)
    Lq = 12
    shapes2 = as_shapes2(T)
    S = sum(T)
    value = rs.standard_normal((N, S, M, D))
    x = rs.uniform(0.05, 0.95, (N, Lq, M, L, P))
    x = keep_off_knots(x, T)
    Tl = np.asarray(T, dtype=np.float64)[None, None, :, None]
    x[:, 0] = 0.0
    x[:, 1] = 1.0
    x[:, 2] = 0.5 / Tl                    # pixel 0 exactly: border gate closed
    x[:, 3] = 1.0 - 0.5 / Tl              # pixel T-1 exactly
    x[:, 4] = (2.0 + 1e-3 + 0.5) / Tl     # just above knot 2
    x[:, 5] = (2.0 - 1e-3 + 0.5) / Tl     # just below knot 2
    x[:, 6] = -1.0
    x[:, 7] = 2.0
    x[:, 8] = (1.25 + 0.5) / Tl           # every sample of the query on the same row pair
    x[:, 9] = -0.3 / Tl                   # pixel in (-1, 0): ZEROS keeps half a sample, BORDER clamps
    x[:, 10] = 1.0 + 0.2 / Tl             # pixel in (T-1, T)
    loc = np.stack([x, np.full_like(x, 0.5)], -1)
    attn = _softmax_lp(rs.standard_normal((N, Lq, M, L, P)))
    onehot = np.zeros((L * P,))
    onehot[5] = 1.0
    attn[:, 11] = onehot.reshape(L, P)
    grad_out = rs.standard_normal((N, Lq, M * D))
    return dict(value=value, shapes=shapes2, lsi=level_start_index(shapes2), loc=loc, attn=attn,
                grad_out=grad_out)


# name -> (kwargs for make_core_inputs) ; the golden generator and the tests iterate this table
CORE_CASES = {
    # BASELINE.json configs[0]: YouCook2 decoder cross-attention shape
    "yc2_dec": dict(seed=0, N=1, M=8, D=64, shapes=level_lengths(200), Lq=100, P=4),
    # ActivityNet shape with its odd level (T=13), batch 2, encoder self-attention (Lq = S)
    "anet_enc": dict(seed=1, N=2, M=8, D=64, shapes=level_lengths(100), Lq=188, P=4),
    "p8_d32": dict(seed=2, N=1, M=4, D=32, shapes=level_lengths(40), Lq=33, P=8),
    "p16_d128_l1": dict(seed=3, N=1, M=2, D=128, shapes=[50], Lq=17, P=16),
    "uniform01": dict(seed=4, N=3, M=2, D=64, shapes=level_lengths(64), Lq=21, P=4, lo=0.0, hi=1.0),
    "wide_range": dict(seed=5, N=1, M=2, D=16, shapes=level_lengths(30), Lq=40, P=4, lo=-1.0, hi=2.0),
    # odd head dims exercise the generic (non-vectorised) kernels, like pdvc/ops/test.py:85-86
    "odd_d30": dict(seed=6, N=1, M=2, D=30, shapes=level_lengths(20), Lq=11, P=2),
    "odd_d71": dict(seed=8, N=1, M=1, D=71, shapes=[9, 5], Lq=6, P=3),
    # 2-D problems: the reference's own test shape (pdvc/ops/test.py:21-25) and a ragged one
    "ref_test_2d": dict(seed=3, N=1, M=2, D=2, shapes=[(6, 4), (3, 2)], Lq=2, P=2, lo=0.0, hi=1.0, two_d=True),
    "ragged_2d": dict(seed=9, N=2, M=3, D=5, shapes=[(5, 7), (3, 4), (1, 2)], Lq=6, P=3, lo=-0.2, hi=1.2, two_d=True),
    "vec_2d": dict(seed=10, N=1, M=2, D=64, shapes=[(4, 9), (2, 5)], Lq=10, P=4, lo=-0.1, hi=1.1, two_d=True),
    # caption-head variant: one head of 512 channels, few queries
    "cap_m1_d512": dict(seed=11, N=1, M=1, D=512, shapes=level_lengths(200), Lq=6, P=4),
}


def core_inputs(name):
    if name == "edges":
        return make_edge_inputs()
    return make_core_inputs(**CORE_CASES[name])


ALL_CORE_CASES = list(CORE_CASES) + ["edges"]


def checksum(d):
    """Order-sensitive fp64 checksum of a dict of arrays (to detect input drift)."""
    acc = 0.0
    for k in sorted(d):
        a = np.asarray(d[k], dtype=np.float64).ravel()
        acc += float(np.dot(a, np.cos(np.arange(a.size) * 0.37 + 1.0)))
    return acc


# ---------------------------------------------------------------------------------------------------------------
# whole-model (PDVC) fixtures: parameters are filled from a per-name seeded formula instead of being stored, so the
# reference model (tests/golden/make_golden_pdvc.py, build container) and this package's model (tests) hold identical
# weights without an 80 MB fixture.
# ---------------------------------------------------------------------------------------------------------------
def canonical_param_name(key):
    """Aliases of one shared tensor map to one name: the decoder's refinement heads are the model's bbox heads
    (pdvc/pdvc.py:93-99) and, with share_caption_head, every caption_head.<i> is caption_head.0 (pdvc.py:80-82)."""
    import re
    key = key.replace("transformer.decoder.bbox_head.", "bbox_head.")
    return re.sub(r"^caption_head\.\d+\.", "caption_head.0.", key)


def fill_parameters(model, share_caption_head=True):
    """Overwrite every parameter/buffer of `model` (reference PDVC or ours: same state_dict keys) in place."""
    import re
    import zlib
    import torch
    with torch.no_grad():
        for key, t in model.state_dict().items():
            if not t.dtype.is_floating_point:
                continue
            name = canonical_param_name(key) if share_caption_head else key.replace("transformer.decoder.bbox_head.", "bbox_head.")
            rs = np.random.RandomState(zlib.crc32(name.encode()) & 0x7FFFFFFF)
            z = rs.standard_normal(tuple(t.shape))
            if t.dim() >= 2:
                new = z * (0.5 / np.sqrt(float(np.prod(t.shape[1:]))))
            elif re.search(r"(norm\d*\.weight|input_proj\.\d+\.1\.weight)$", name):
                new = 1.0 + 0.05 * z
            elif name.endswith("sampling_offsets.bias"):
                new = t.detach().double().cpu().numpy() + 0.1 * z
            else:
                new = 0.05 * z
            t.copy_(torch.from_numpy(np.ascontiguousarray(new)).to(t.dtype))


def pdvc_batch(seed, T, feature_dim, n_gt, cap_len, vocab_size, dtype="float64"):
    """Synthetic single-video batch as numpy arrays."""
    rs = np.random.RandomState(seed)
    words = cap_len + 4
    centre, length = rs.uniform(0.2, 0.8, n_gt), rs.uniform(0.05, 0.25, n_gt)
    caps = np.zeros((n_gt, words), dtype=np.int64)
    caps[:, 1:1 + cap_len] = rs.randint(1, vocab_size, (n_gt, cap_len))
    # ragged sentence lengths: some captions end early
    for g in range(n_gt):
        caps[g, 1 + cap_len - (g % 3):] = 0
    mask = np.zeros((n_gt, words))
    for g in range(n_gt):
        mask[g, :cap_len - (g % 3) + 2] = 1.0
    valid = np.ones((1, T), dtype=bool)
    valid[0, T - T // 8:] = False            # padded tail: exercises the padding mask and valid ratios
    return dict(video_tensor=rs.standard_normal((1, T, feature_dim)).astype(dtype), video_mask=valid,
                video_length=np.array([[float(T), 117.6, float(n_gt)]], dtype=dtype),
                boxes=np.stack([centre, length], -1).astype(dtype), labels=np.zeros(n_gt, dtype=np.int64),
                cap_tensor=caps, cap_mask=mask.astype(dtype))
