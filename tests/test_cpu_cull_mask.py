"""The K3 footprint cull (csrc/geometry.cu: tile_cull_mask) must be CONSERVATIVE: a warp footprint may only be dropped
when none of its 8x4 pixels can reach the splat's skip threshold.  The GPU suite proves this indirectly (bit-exact
n_contrib / final_T / images on every scene); this CPU test restates the shared-line formula in float32 numpy and
checks it against a float64 brute force over the 32 pixels on adversarial inputs the scenes do not contain
(needle-like conics, centres hundreds of pixels away, thresholds from tight to loose).  No reference counterpart: the
reference has no footprint cull (renderCUDA, forward.cu:261-374, evaluates every pixel of the tile)."""
import numpy as np

f32 = np.float32


def tile_cull_mask_f32(cx, cy, A, B, C, thr, tx, ty):
    """Vectorised float32 restatement of tile_cull_mask (no FMA: every product and sum rounds - a little worse than
    the kernel, which is what a conservativeness test wants).  Returns uint32 masks, bit w = warp footprint w kept."""
    cx, cy, A, B, C, thr, tx, ty = [np.asarray(v, dtype=f32) for v in (cx, cy, A, B, C, thr, tx, ty)]
    n = cx.shape[0]
    bad = ~(A > 0) | ~(C > 0) | ~(B == B)
    with np.errstate(all="ignore"):
        invC, invA = f32(1) / C, f32(1) / A
        nBiC, nBiA = -B * invC, -B * invA
        Dx, Dy = B * nBiC + A, B * nBiA + C
        X0, Y0 = cx - tx, cy - ty
        ax = np.maximum(np.abs(X0), np.abs(X0 - f32(15)))
        ay = np.maximum(np.abs(Y0), np.abs(Y0 - f32(15)))
        qcut = f32(2) * (f32(2e-3) - thr) + f32(1e-6) * (A * ax * ax + C * ay * ay)
        ysl, basel = [], []
        for l in range(4):
            X = X0 - f32((l >> 1) * 8 + (l & 1) * 7)
            ysl.append(nBiC * X)
            basel.append(Dx * X * X)
        inx = [(X0 - f32(7) <= 0) & (X0 >= 0), (X0 - f32(15) <= 0) & (X0 - f32(8) >= 0)]
        mask = np.zeros(n, dtype=np.uint32)
        for k in range(4):
            y1 = Y0 - f32(4 * k)
            y0 = y1 - f32(3)
            qv = []
            for l in range(4):
                e = np.minimum(np.maximum(ysl[l], y0), y1) - ysl[l]
                qv.append(C * e * e + basel[l])
            qh = [np.full(n, 3.0e38, dtype=f32), np.full(n, 3.0e38, dtype=f32)]
            for Y in (y1, y0):
                xs, base = nBiA * Y, Dy * Y * Y
                for b in range(2):
                    x1 = X0 - f32(8 * b)
                    x0 = x1 - f32(7)
                    e = np.minimum(np.maximum(xs, x0), x1) - xs
                    qh[b] = np.fmin(qh[b], A * e * e + base)
            iny = (y0 <= 0) & (y1 >= 0)
            q0 = np.fmin(np.fmin(qv[0], qv[1]), qh[0])
            q1 = np.fmin(np.fmin(qv[2], qv[3]), qh[1])
            keep0 = (inx[0] & iny) | ~(q0 * f32(1 - 2e-5) > qcut)
            keep1 = (inx[1] & iny) | ~(q1 * f32(1 - 2e-5) > qcut)
            mask |= keep0.astype(np.uint32) << np.uint32(2 * k)
            mask |= keep1.astype(np.uint32) << np.uint32(2 * k + 1)
    mask[bad] = 0xFF
    return mask


def max_power_per_footprint(cx, cy, A, B, C, tx, ty):
    """float64 brute force: the largest power = -q/2 over the 32 pixels of each of the 8 footprints -> [n, 8]."""
    cx, cy, A, B, C, tx, ty = [np.asarray(v, dtype=np.float64) for v in (cx, cy, A, B, C, tx, ty)]
    out = np.empty((cx.shape[0], 8))
    for w in range(8):
        px = tx[:, None] + (w & 1) * 8 + np.arange(8)[None, :]                # [n, 8]
        py = ty[:, None] + (w >> 1) * 4 + np.arange(4)[None, :]               # [n, 4]
        dx = (cx[:, None] - px)[:, :, None]                                   # [n, 8, 1]
        dy = (cy[:, None] - py)[:, None, :]                                   # [n, 1, 4]
        q = A[:, None, None] * dx * dx + 2.0 * B[:, None, None] * dx * dy + C[:, None, None] * dy * dy
        out[:, w] = (-0.5 * q).reshape(cx.shape[0], -1).max(axis=1)
    return out


def _cases(rng, n):
    # conics from covariance eigen-decompositions: sigma_major in [0.6, 400] px, axis ratio up to 1 : 600 (needles),
    # plus the 0.3 low-pass the preprocess adds to the covariance diagonal (forward.cu:112-113)
    s1 = np.exp(rng.uniform(np.log(0.6), np.log(400.0), n))
    s2 = s1 / np.exp(rng.uniform(0.0, np.log(600.0), n))
    th = rng.uniform(0, np.pi, n)
    c, s = np.cos(th), np.sin(th)
    a = c * c * s1 * s1 + s * s * s2 * s2 + 0.3
    b = c * s * (s1 * s1 - s2 * s2)
    d = s * s * s1 * s1 + c * c * s2 * s2 + 0.3
    det = a * d - b * b
    A, B, C = d / det, -b / det, a / det
    tx = rng.integers(0, 240, n) * 16.0
    ty = rng.integers(0, 135, n) * 16.0
    # centres from inside the tile to ~3.5 sigma_major away (the radius the reference bins with is 3 sigma)
    r = rng.uniform(0, 3.5, n) * s1 * (rng.uniform(0, 1, n) < 0.7) + rng.uniform(-2, 18, n)
    phi = rng.uniform(0, 2 * np.pi, n)
    cx = tx + 8 + r * np.cos(phi)
    cy = ty + 8 + r * np.sin(phi)
    opacity = np.exp(rng.uniform(np.log(1 / 254.0), 0.0, n))
    thr = -np.log(255.0 * opacity) - 0.01
    return cx, cy, A, B, C, thr, tx, ty


def test_cull_mask_never_drops_a_footprint_that_can_contribute():
    rng = np.random.default_rng(20240607)
    n = 400_000
    cx, cy, A, B, C, thr, tx, ty = _cases(rng, n)
    cx32, cy32, A32, B32, C32, thr32 = [v.astype(f32) for v in (cx, cy, A, B, C, thr)]
    mask = tile_cull_mask_f32(cx32, cy32, A32, B32, C32, thr32, tx, ty)
    # ground truth on the float32 inputs the kernel sees, evaluated in float64
    pmax = max_power_per_footprint(cx32, cy32, A32, B32, C32, tx, ty)
    # the blend takes a pixel only if alpha = opacity * exp(power) >= 1/255, i.e. power >= thr + 0.01; leave half of
    # that margin to the float evaluation of power in the blend kernel
    must_keep = (pmax >= thr32.astype(np.float64)[:, None] + 0.005) & (pmax <= 0.0)
    kept = ((mask[:, None] >> np.arange(8, dtype=np.uint32)[None, :]) & 1).astype(bool)
    bad = must_keep & ~kept
    assert not bad.any(), f"{int(bad.sum())} contributing footprints dropped, e.g. case {np.argwhere(bad)[0]}"
    # and it does cull: most footprints whose best pixel is far below the threshold go (not a correctness property -
    # a regression check that the margins have not been inflated into uselessness)
    hopeless = pmax < thr32.astype(np.float64)[:, None] - 1.0
    assert hopeless.sum() > 1000
    assert (kept & hopeless).sum() <= 0.02 * hopeless.sum(), float((kept & hopeless).sum()) / float(hopeless.sum())


def test_cull_mask_degenerate_inputs_keep_everything():
    one = np.ones(6, dtype=f32)
    tx = ty = np.zeros(6, dtype=f32)
    A = np.array([0.0, -1.0, 1.0, 1.0, np.nan, 1.0], dtype=f32)
    B = np.array([0.0, 0.0, np.nan, 0.0, 0.0, 0.0], dtype=f32)
    C = np.array([1.0, 1.0, 1.0, 0.0, 1.0, np.nan], dtype=f32)
    m = tile_cull_mask_f32(500 * one, 500 * one, A, B, C, -3 * one, tx, ty)
    assert (m == 0xFF).all()
    # opacity <= 0 (threshold 1e30: every power skips) culls every footprint the centre is not inside; a NaN opacity
    # (threshold -1e30) keeps all
    m = tile_cull_mask_f32(np.array([100.0, 3.0], dtype=f32), np.array([100.0, 2.0], dtype=f32), one[:2], 0 * one[:2],
                           one[:2], np.array([1e30, 1e30], dtype=f32), tx[:2], ty[:2])
    assert m[0] == 0 and m[1] == 1
    m = tile_cull_mask_f32(np.array([100.0], dtype=f32), np.array([100.0], dtype=f32), one[:1], 0 * one[:1], one[:1],
                           np.array([-1e30], dtype=f32), tx[:1], ty[:1])
    assert m[0] == 0xFF
