"""TEST INFRASTRUCTURE - numpy restatement of the host side of the reference's eval grid.

Only tests/, __graft_entry__.smoke() and bench.py's reference / cpu_baseline legs may import this module; the
product path (r3dgw_b200.envmap_sh) never does.

  basis_matrix / solid_angle_rows / project   utils/sh_additional_utils.py:95-109, 159-206 and
                                              utils/spherical_harmonics.py:43-78, 114-126, 148-216
      PINNED: tests/golden/envmap_sh.npz holds coefficients produced by the reference's own
      get_coefficients_from_image (imported unmodified by tests/golden/make_envmap_golden.py).
  scale_bright                                eval_with_gt_envmaps.py:42-49 (process_environment_map_image)
  rotation_euler / sh_rotation_matrix / rotate_sh
      The reference calls spaudiopy.sph.rotate_sh(F, yaw, pitch, roll, 'real') (eval_with_gt_envmaps.py:97-98,
      relit_novel_view.py:72-74).  spaudiopy (environment.yml:239, ==0.1.6) is a third-party dependency that is
      absent from /root/reference and from this image, so this is a restatement of its published algorithm: the
      real-SH rotation recursion of Ivanic & Ruedenberg (J. Phys. Chem. 1996, 100, 6342; errata 1998, 102, 9099)
      seeded with the Cartesian z-y-x Euler matrix re-ordered to (y, z, x), rows of coefficients multiplied by the
      transposed block-diagonal matrix.  PARITY UNPINNED against spaudiopy itself; tests anchor it on the
      properties the algorithm must have (orthogonality, D(R1 R2) = D(R1) D(R2), band 1 = permuted R, and
      Y(R^T d) = D^T Y(d) for an independently evaluated real ACN basis).
"""
import math

import numpy as np


# ---- projection -------------------------------------------------------------------------------------------------
def _legendre(l, m, x):
    """utils/spherical_harmonics.py:43-78 (associated Legendre with the Condon-Shortley sign)."""
    pmm = np.ones_like(x)
    if m > 0:
        s = np.sqrt((1.0 - x) * (1.0 + x))
        f = 1.0
        for _ in range(m):
            pmm = pmm * (-f) * s
            f += 2.0
    if l == m:
        return pmm
    pm1 = x * (2.0 * m + 1.0) * pmm
    if l == m + 1:
        return pm1
    out = pm1
    for ll in range(m + 2, l + 1):
        out = ((2.0 * ll - 1.0) * x * pm1 - (ll + m - 1.0) * pmm) / (ll - m)
        pmm, pm1 = pm1, out
    return out


def _norm(l, m):
    """utils/spherical_harmonics.py:114-126."""
    return math.sqrt(((2 * l + 1) * math.factorial(l - m)) / (4 * math.pi * math.factorial(l + m)))


def basis_matrix(width, l_max):
    """sh_additional_utils.py:159-170 -> [H, W, (l_max+1)^2]; theta = y*pi/H (no half-pixel offset), phi = x*2pi/W."""
    height = int(width / 2)
    theta = (np.arange(height, dtype=np.float64) / (float(height) / np.pi)).reshape(height, 1)
    phi = np.arange(width, dtype=np.float64) / (float(width) / (np.pi * 2))
    ct = np.cos(theta)
    Y = np.zeros((height, width, (l_max + 1) ** 2))
    for l in range(l_max + 1):
        for m in range(-l, l + 1):
            a = abs(m)
            leg = _legendre(l, a, ct)
            if m == 0:
                v = _norm(l, 0) * leg * np.ones_like(phi)
            elif m > 0:
                v = math.sqrt(2.0) * _norm(l, a) * np.cos(m * phi) * leg
            else:
                v = math.sqrt(2.0) * _norm(l, a) * np.sin(a * phi) * leg
            Y[:, :, l * l + l + m] = v
    return Y


def solid_angle_rows(width):
    """sh_additional_utils.py:95-105, one value per row."""
    height = int(width / 2)
    y = np.arange(height)
    theta = (1.0 - ((y + 0.5) / height)) * np.pi
    half = (np.pi / height) / 2.0
    return ((np.pi * 2) / width) * (np.cos(theta - half) - np.cos(theta + half))


def project(ibl, l_max=2):
    """sh_additional_utils.py:176-206 for an image that needs no resize: [H,W,>=3] -> f64 [(l_max+1)^2, 3]."""
    ibl = np.asarray(ibl)
    H, W = ibl.shape[:2]
    Y = basis_matrix(W, l_max)
    sa = np.repeat(solid_angle_rows(W)[:, None], W, axis=1)
    n = (l_max + 1) ** 2
    out = np.zeros((n, 3))
    for i in range(n):
        for c in range(3):
            out[i, c] = np.sum(ibl[:, :, c] * Y[:, :, i] * sa)
    return out


def scale_bright(img_u8, scale_high, threshold):
    """eval_with_gt_envmaps.py:44-47: u8 -> f32/255, entries above the threshold multiplied by scale_high."""
    img = img_u8.astype(np.float32) / np.float32(255)
    img = img.copy()
    img[img > threshold] *= scale_high
    return img


# ---- rotation (spaudiopy 0.1.6 sph.rotate_sh restated; see the header) -------------------------------------------
def rotation_euler(yaw, pitch, roll):
    """z-y-x Euler matrix Rz(yaw) Ry(pitch) Rx(roll)."""
    cz, sz = math.cos(yaw), math.sin(yaw)
    cy, sy = math.cos(pitch), math.sin(pitch)
    cx, sx = math.cos(roll), math.sin(roll)
    Rz = np.array([[cz, -sz, 0.], [sz, cz, 0.], [0., 0., 1.]])
    Ry = np.array([[cy, 0., sy], [0., 1., 0.], [-sy, 0., cy]])
    Rx = np.array([[1., 0., 0.], [0., cx, -sx], [0., sx, cx]])
    return Rz @ Ry @ Rx


def _ir_p(i, l, a, b, r1, prev):
    """Ivanic-Ruedenberg helper P_i(l; a, b): r1 is the band-1 block indexed [-1..1]+1, prev the band l-1 block."""
    ri1, rim1, ri0 = r1[i + 1, 2], r1[i + 1, 0], r1[i + 1, 1]
    if b == -l:
        return ri1 * prev[a + l - 1, 0] + rim1 * prev[a + l - 1, 2 * l - 2]
    if b == l:
        return ri1 * prev[a + l - 1, 2 * l - 2] - rim1 * prev[a + l - 1, 0]
    return ri0 * prev[a + l - 1, b + l - 1]


def _ir_u(l, m, n, r1, prev):
    return _ir_p(0, l, m, n, r1, prev)


def _ir_v(l, m, n, r1, prev):
    if m == 0:
        return _ir_p(1, l, 1, n, r1, prev) + _ir_p(-1, l, -1, n, r1, prev)
    if m > 0:
        d = 1.0 if m == 1 else 0.0
        return _ir_p(1, l, m - 1, n, r1, prev) * math.sqrt(1 + d) - _ir_p(-1, l, -m + 1, n, r1, prev) * (1 - d)
    d = 1.0 if m == -1 else 0.0
    return _ir_p(1, l, m + 1, n, r1, prev) * (1 - d) + _ir_p(-1, l, -m - 1, n, r1, prev) * math.sqrt(1 + d)


def _ir_w(l, m, n, r1, prev):
    if m > 0:
        return _ir_p(1, l, m + 1, n, r1, prev) + _ir_p(-1, l, -m - 1, n, r1, prev)
    return _ir_p(1, l, m - 1, n, r1, prev) - _ir_p(-1, l, -m + 1, n, r1, prev)


def sh_rotation_blocks(l_max, yaw, pitch, roll):
    """Per-band real-SH rotation blocks [(2l+1) x (2l+1)], l = 0..l_max."""
    R = rotation_euler(yaw, pitch, roll)
    blocks = [np.ones((1, 1))]
    if l_max == 0:
        return blocks
    r1 = R[np.ix_([1, 2, 0], [1, 2, 0])]
    blocks.append(r1)
    prev = r1
    for l in range(2, l_max + 1):
        cur = np.zeros((2 * l + 1, 2 * l + 1))
        for m in range(-l, l + 1):
            for n in range(-l, l + 1):
                d = 1.0 if m == 0 else 0.0
                denom = (2 * l) * (2 * l - 1) if abs(n) == l else (l * l - n * n)
                u = math.sqrt((l * l - m * m) / denom)
                v = math.sqrt((1 + d) * (l + abs(m) - 1) * (l + abs(m)) / denom) * (1 - 2 * d) * 0.5
                w = math.sqrt((l - abs(m) - 1) * (l - abs(m)) / denom) * (1 - d) * (-0.5)
                acc = 0.0
                if u != 0:
                    acc += u * _ir_u(l, m, n, r1, prev)
                if v != 0:
                    acc += v * _ir_v(l, m, n, r1, prev)
                if w != 0:
                    acc += w * _ir_w(l, m, n, r1, prev)
                cur[m + l, n + l] = acc
        blocks.append(cur)
        prev = cur
    return blocks


def sh_rotation_matrix(l_max, yaw, pitch, roll):
    blocks = sh_rotation_blocks(l_max, yaw, pitch, roll)
    n = (l_max + 1) ** 2
    D = np.zeros((n, n))
    for l, b in enumerate(blocks):
        D[l * l:(l + 1) ** 2, l * l:(l + 1) ** 2] = b
    return D


def rotate_sh(F, yaw, pitch, roll):
    """rows of coefficients [..., (l_max+1)^2] -> rotated rows (F @ D^T)."""
    F = np.asarray(F, dtype=np.float64)
    l_max = int(round(math.sqrt(F.shape[-1]))) - 1
    return F @ sh_rotation_matrix(l_max, float(yaw), float(pitch), float(roll)).T


def eval_grid_rotations(coeffs, init_zyx, angles):
    """eval_with_gt_envmaps.py:95-100 for every sun angle: coeffs [n,3] f64, init_zyx = (init_rot_z, init_rot_y,
    init_rot_x), angles [B,3] with rows (a0, a1, a2) used as rotate_sh(., a2, a0, a1) -> f32 [B, n, 3]."""
    base = rotate_sh(np.asarray(coeffs).T, *init_zyx)
    out = []
    for a in np.asarray(angles, dtype=np.float64):
        out.append(rotate_sh(base, a[2], a[0], a[1]).T.astype(np.float32))
    return np.stack(out)


def real_sh_acn(l_max, dirs):
    """Independent closed-form real orthonormal SH in ACN order WITHOUT the Condon-Shortley sign (the basis the
    rotation recursion is defined for), from unit vectors [K,3]; used only by the rotation property tests."""
    x, y, z = dirs[:, 0], dirs[:, 1], dirs[:, 2]
    phi = np.arctan2(y, x)
    out = np.zeros((dirs.shape[0], (l_max + 1) ** 2))
    for l in range(l_max + 1):
        for m in range(-l, l + 1):
            a = abs(m)
            leg = _legendre(l, a, z) * ((-1.0) ** a)          # strip the Condon-Shortley sign
            k = _norm(l, a)
            if m == 0:
                v = k * leg
            elif m > 0:
                v = math.sqrt(2.0) * k * np.cos(a * phi) * leg
            else:
                v = math.sqrt(2.0) * k * np.sin(a * phi) * leg
            out[:, l * l + l + m] = v
    return out


# ---- cv2.resize(img, (w, h), interpolation=cv2.INTER_CUBIC) restated (opencv-python==4.10.0.84, environment.yml:210)
def _cubic_weights(x):
    A = np.float32(-0.75)
    x = x.astype(np.float32)
    one = np.float32(1)
    w0 = ((A * (x + one) - np.float32(5) * A) * (x + one) + np.float32(8) * A) * (x + one) - np.float32(4) * A
    w1 = ((A + np.float32(2)) * x - (A + np.float32(3))) * x * x + one
    w2 = ((A + np.float32(2)) * (one - x) - (A + np.float32(3))) * (one - x) * (one - x) + one
    w3 = one - w0 - w1 - w2
    return np.stack([w0, w1, w2, w3]).astype(np.float32)


def _cubic_taps(n_src, n_dst):
    scale = float(n_src) / n_dst
    f = ((np.arange(n_dst) + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int64)
    frac = (f - s.astype(np.float32)).astype(np.float32)
    idx = np.clip(s[None, :] + np.arange(-1, 3)[:, None], 0, n_src - 1)
    return idx, _cubic_weights(frac)


def resize_cubic(img, width, height):
    """OpenCV's generic f32 cubic resize: 4 clamped taps per axis, A = -0.75, horizontal pass then vertical pass,
    f32 throughout, products accumulated left to right."""
    img = np.asarray(img, dtype=np.float32)[:, :, :3]
    xi, xw = _cubic_taps(img.shape[1], width)
    yi, yw = _cubic_taps(img.shape[0], height)
    h = np.zeros((img.shape[0], width, 3), np.float32)
    for k in range(4):
        h = (h + img[:, xi[k], :] * xw[k][None, :, None]).astype(np.float32) if k else \
            (img[:, xi[k], :] * xw[k][None, :, None]).astype(np.float32)
    v = None
    for k in range(4):
        term = (h[yi[k], :, :] * yw[k][:, None, None]).astype(np.float32)
        v = term if v is None else (v + term).astype(np.float32)
    return v


def coefficients_from_image(ibl, l_max=2, resize_width=None):
    """sh_additional_utils.py:176-206 including the resize rule of :178-184."""
    ibl = np.asarray(ibl)
    if resize_width is not None:
        ibl = resize_cubic(ibl, resize_width, int(resize_width / 2))
    elif ibl.shape[1] > 1000:
        ibl = resize_cubic(ibl, 1000, 500)
    return project(ibl, l_max)


def synth_envmap(w, phase=0.0):
    """Deterministic HDR-like lat-long test image f32 [w/2, w, 3] (no RNG state: tests regenerate the large inputs
    whose reference outputs are stored in tests/golden/envmap_sh.npz)."""
    h = w // 2
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    sky = 0.3 + 0.5 * (1 - yy / max(h, 1))
    sun = 8.0 * np.exp(-(((xx - (0.3 + phase) * w) / (0.03 * w + 1)) ** 2 + ((yy - 0.25 * h) / (0.03 * w + 1)) ** 2))
    noise = np.modf(np.abs(np.sin(xx * 12.9898 + yy * 78.233 + phase) * 43758.5453))[0]
    img = (sky + sun)[..., None] * np.array([1.0, 0.9, 0.8]) + 0.1 * noise[..., None] * np.array([1.0, 0.5, 0.25])
    return img.astype(np.float32)
