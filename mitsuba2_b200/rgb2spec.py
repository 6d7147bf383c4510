"""sRGB -> spectrum upsampling model of the spectral variants (host side).

The reference converts every RGB colour it meets in a spectral variant into three coefficients of the
Jakob & Hanika (2019) sigmoid-polynomial model, ``srgb_model_fetch`` (src/librender/srgb.cpp:15-42), by trilinear
lookup in a 3 x 64^3 table ``data/srgb.coeff`` that its build generates by running ``rgb2spec_opt 64``
(ext/rgb2spec/CMakeLists.txt:51).  There is no such file in a checkout, so this module restates both halves:

* :func:`build_table` -- the optimisation of ext/rgb2spec/rgb2spec_opt.cpp:113-270, 288-420 (Gauss-Newton on the
  CIELAB residual under D65, finite-difference Jacobian, walking outwards in brightness from ``res/5`` with the
  previous solution as initial guess), vectorised over the (axis, y, x) lanes (torch, float64; on the GPU when present);
* :func:`fetch` -- ``rgb2spec_fetch`` (ext/rgb2spec/rgb2spec.c:75-118) incl. the special cases of srgb.cpp:34-37.

The table is cached next to this file (``_srgb_coeff_<res>.npz``, git-ignored); building resolution 64 takes a few
seconds on a GPU and about two minutes on host cores.
The device evaluates the coefficients (``srgb_model_eval``, csrc/spectrum.h); it never sees RGB in spectral mode.
"""
import os

import numpy as np

from ._cie import CIE_D65, CIE_MAX, CIE_MIN, CIE_SAMPLES, CIE_X, CIE_Y, CIE_Z

XYZ_TO_SRGB = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]])
SRGB_TO_XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]])
FINE = (CIE_SAMPLES - 1) * 3 + 1
EPS = 1e-4
_tables = {}


def _cie_interp(data, lam):
    x = (lam - CIE_MIN) * ((CIE_SAMPLES - 1) / (CIE_MAX - CIE_MIN))
    off = np.clip(x.astype(np.int64), 0, CIE_SAMPLES - 2)
    w = x - off
    data = np.asarray(data, dtype=np.float64)
    return (1.0 - w) * data[off] + w * data[off + 1]


def _init_tables():
    """init_tables(SRGB): Simpson 3/8 quadrature weights folded into spectrum->rgb rows and the whitepoint."""
    h = (CIE_MAX - CIE_MIN) / (FINE - 1)
    i = np.arange(FINE)
    lam = CIE_MIN + i * h
    xyz = np.stack([_cie_interp(CIE_X, lam), _cie_interp(CIE_Y, lam), _cie_interp(CIE_Z, lam)])          # 3 x FINE
    illum = _cie_interp(np.asarray(CIE_D65) / 10566.864005283874576, lam)
    weight = np.full(FINE, 3.0 / 8.0 * h)
    inner = (i > 0) & (i < FINE - 1)
    weight[inner & ((i - 1) % 3 == 2)] *= 2.0
    weight[inner & ((i - 1) % 3 != 2)] *= 3.0
    rgb_tbl = XYZ_TO_SRGB @ (xyz * illum * weight)                                                        # 3 x FINE
    white = (xyz * illum * weight).sum(axis=1)
    return (lam - CIE_MIN) / (CIE_MAX - CIE_MIN), rgb_tbl, white


_LAM01, _RGB_TBL, _WHITE = _init_tables()


def _cie_lab(rgb):
    xyz = rgb @ SRGB_TO_XYZ.T
    t = xyz / _WHITE
    delta = 6.0 / 29.0
    f = np.where(t > delta ** 3, np.cbrt(np.maximum(t, 0.0)), t / (delta * delta * 3.0) + 4.0 / 29.0)
    return np.stack([116.0 * f[..., 1] - 16.0, 500.0 * (f[..., 0] - f[..., 1]), 200.0 * (f[..., 1] - f[..., 2])], axis=-1)


def _residual(coeffs, lab_target):
    """eval_residual for a batch: coeffs (..., 3) on the normalised wavelength axis."""
    x = (coeffs[..., 0:1] * _LAM01 + coeffs[..., 1:2]) * _LAM01 + coeffs[..., 2:3]
    s = 0.5 * x / np.sqrt(1.0 + x * x) + 0.5
    return lab_target - _cie_lab(s @ _RGB_TBL.T)


def gauss_newton(rgb, coeffs, iterations=15):
    """gauss_newton() of rgb2spec_opt.cpp:233-270 for a batch of lanes; updates and returns ``coeffs``."""
    lab = _cie_lab(rgb)
    active = np.ones(rgb.shape[0], dtype=bool)
    for _ in range(iterations):
        idx = np.nonzero(active)[0]
        if idx.size == 0:
            break
        c = coeffs[idx]
        r = _residual(c, lab[idx])
        jac = np.empty((idx.size, 3, 3))
        for k in range(3):
            d = np.zeros(3)
            d[k] = EPS
            jac[:, :, k] = (_residual(c + d, lab[idx]) - _residual(c - d, lab[idx])) / (2 * EPS)
        det = np.linalg.det(jac)
        ok = np.abs(det) > 1e-300
        step = np.zeros_like(c)
        if ok.any():
            step[ok] = np.linalg.solve(jac[ok], r[ok][..., None])[..., 0]
        c = c - step
        mx = c.max(axis=1)
        big = mx > 200
        c[big] *= (200.0 / mx[big])[:, None]
        coeffs[idx] = c
        active[idx[(r * r).sum(axis=1) < 1e-6]] = False
    return coeffs


def _smoothstep(x):
    return x * x * (3.0 - 2.0 * x)


def _gauss_newton_torch(rgb, coeffs, lam01, rgb_tbl_t, to_xyz_t, white, iterations=15):
    """The same iteration as :func:`gauss_newton` with torch tensors (float64) and fixed shapes: converged lanes are
    frozen with a mask instead of being compacted, the 3x3 systems are solved with Cramer's rule."""
    import torch
    delta = 6.0 / 29.0

    def lab(c):
        t = (c @ to_xyz_t) / white
        f = torch.where(t > delta ** 3, torch.sign(t) * torch.abs(t).pow(1.0 / 3.0), t / (delta * delta * 3.0) + 4.0 / 29.0)
        return torch.stack([116.0 * f[:, 1] - 16.0, 500.0 * (f[:, 0] - f[:, 1]), 200.0 * (f[:, 1] - f[:, 2])], dim=1)

    def residual(c, target):
        x = (c[:, 0:1] * lam01 + c[:, 1:2]) * lam01 + c[:, 2:3]
        s = 0.5 * x / torch.sqrt(1.0 + x * x) + 0.5
        return target - lab(s @ rgb_tbl_t)

    target = lab(rgb)
    active = torch.ones(rgb.shape[0], dtype=torch.bool, device=rgb.device)
    eye = torch.eye(3, dtype=rgb.dtype, device=rgb.device) * EPS
    for _ in range(iterations):
        r = residual(coeffs, target)
        cols = [(residual(coeffs + eye[k], target) - residual(coeffs - eye[k], target)) / (2 * EPS) for k in range(3)]
        a, b, c = cols                                     # columns of the Jacobian
        det = (a * torch.linalg.cross(b, c)).sum(dim=1)
        ok = det.abs() > 1e-300
        inv = 1.0 / torch.where(ok, det, torch.ones_like(det))
        step = torch.stack([(r * torch.linalg.cross(b, c)).sum(dim=1), (a * torch.linalg.cross(r, c)).sum(dim=1),
                            (a * torch.linalg.cross(b, r)).sum(dim=1)], dim=1) * inv[:, None]
        new = coeffs - torch.where(ok[:, None], step, torch.zeros_like(step))
        mx = new.max(dim=1).values
        new = torch.where((mx > 200)[:, None], new * (200.0 / mx)[:, None], new)
        coeffs = torch.where(active[:, None], new, coeffs)
        active = active & ~((r * r).sum(dim=1) < 1e-6)
        if not bool(active.any()):
            break
    return coeffs


def build_table(res=64, progress=None, device=None):
    """main() of rgb2spec_opt.cpp:288-420 -> (scale[res] float32, data[3, res, res, res, 3] float32).  Runs in float64 on
    the GPU when there is one (a second or two), otherwise on the host cores (about two minutes at resolution 64)."""
    if device is None:
        import torch
        device = "cuda" if torch.cuda.is_available() else "numpy"
    if device == "numpy":
        return _build_table_numpy(res, progress)
    import torch
    kw = dict(dtype=torch.float64, device=device)
    lam01, rgb_tbl_t = torch.tensor(_LAM01, **kw), torch.tensor(_RGB_TBL.T.copy(), **kw)
    to_xyz_t, white = torch.tensor(SRGB_TO_XYZ.T.copy(), **kw), torch.tensor(_WHITE, **kw)
    scale = _smoothstep(_smoothstep(np.arange(res) / float(res - 1))).astype(np.float32)
    out = np.zeros((3, res, res, res, 3), dtype=np.float32)
    jj, ii = np.meshgrid(np.arange(res), np.arange(res), indexing="ij")
    y = torch.tensor((jj / float(res - 1)).reshape(-1), **kw)
    x = torch.tensor((ii / float(res - 1)).reshape(-1), **kw)
    c0, c1 = 360.0, 1.0 / (830.0 - 360.0)
    start = res // 5
    for l in range(3):
        for ks in (range(start, res), range(start, -1, -1)):
            coeffs = torch.zeros((res * res, 3), **kw)
            for k in ks:
                b = float(scale[k])
                cols = [None, None, None]
                cols[l], cols[(l + 1) % 3], cols[(l + 2) % 3] = torch.full_like(x, b), x * b, y * b
                coeffs = _gauss_newton_torch(torch.stack(cols, dim=1), coeffs, lam01, rgb_tbl_t, to_xyz_t, white)
                A, B, Cc = coeffs[:, 0], coeffs[:, 1], coeffs[:, 2]
                cell = torch.stack([A * c1 * c1, B * c1 - 2 * A * c0 * c1 * c1, Cc - B * c0 * c1 + A * (c0 * c1) ** 2], dim=1)
                out[l, k] = cell.reshape(res, res, 3).to(torch.float32).cpu().numpy()
                if progress:
                    progress(l, k)
    return scale, out


def _build_table_numpy(res, progress=None):
    scale = _smoothstep(_smoothstep(np.arange(res) / float(res - 1))).astype(np.float32)
    out = np.zeros((3, res, res, res, 3), dtype=np.float32)
    jj, ii = np.meshgrid(np.arange(res), np.arange(res), indexing="ij")
    y = (jj / float(res - 1)).reshape(-1)
    x = (ii / float(res - 1)).reshape(-1)
    c0, c1 = 360.0, 1.0 / (830.0 - 360.0)
    start = res // 5
    for l in range(3):
        for ks in (range(start, res), range(start, -1, -1)):
            coeffs = np.zeros((res * res, 3))
            for k in ks:
                b = float(scale[k])
                rgb = np.empty((res * res, 3))
                rgb[:, l], rgb[:, (l + 1) % 3], rgb[:, (l + 2) % 3] = b, x * b, y * b
                coeffs = gauss_newton(rgb, coeffs)
                A, B, Cc = coeffs[:, 0], coeffs[:, 1], coeffs[:, 2]
                cell = np.stack([A * c1 * c1, B * c1 - 2 * A * c0 * c1 * c1, Cc - B * c0 * c1 + A * (c0 * c1) ** 2], axis=1)
                out[l, k] = cell.reshape(res, res, 3).astype(np.float32)
                if progress:
                    progress(l, k)
    return scale, out


def table(res=64):
    """The cached model table (built on first use)."""
    if res not in _tables:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_srgb_coeff_%d.npz" % res)
        if os.path.exists(path):
            z = np.load(path)
            _tables[res] = (z["scale"], z["data"])
        else:
            scale, data = build_table(res)
            try:
                np.savez(path, scale=scale, data=data)
            except OSError:
                pass
            _tables[res] = (scale, data)
    return _tables[res]


def _find_interval(values, x):
    """rgb2spec_find_interval (rgb2spec.c:54-72): last index with values[i] <= x, clamped to [0, n-2]."""
    return np.clip(np.searchsorted(values, x, side="right") - 1, 0, len(values) - 2)


def fetch(rgb, res=64, model=None):
    """``srgb_model_fetch`` for an array of colours (..., 3) in [0, 1] -> coefficients (..., 3) float32."""
    scale, data = model if model is not None else table(res)
    res = len(scale)
    rgb_in = np.asarray(rgb, dtype=np.float32)
    flat = np.clip(rgb_in.reshape(-1, 3), 0.0, 1.0).astype(np.float32)
    n = flat.shape[0]
    i = np.zeros(n, dtype=np.int64)
    for j in (1, 2):
        i = np.where(flat[:, j] >= flat[np.arange(n), i], j, i)
    z = flat[np.arange(n), i]
    with np.errstate(divide="ignore", invalid="ignore"):
        sc = np.float32(res - 1) / z
        x = flat[np.arange(n), (i + 1) % 3] * sc
        y = flat[np.arange(n), (i + 2) % 3] * sc
    x = np.nan_to_num(x, nan=0.0, posinf=0.0).astype(np.float32)
    y = np.nan_to_num(y, nan=0.0, posinf=0.0).astype(np.float32)
    xi = np.minimum(x.astype(np.int64), res - 2)
    yi = np.minimum(y.astype(np.int64), res - 2)
    zi = _find_interval(scale, z)
    x1 = (x - xi).astype(np.float32)[:, None]
    y1 = (y - yi).astype(np.float32)[:, None]
    z1 = ((z - scale[zi]) / (scale[zi + 1] - scale[zi])).astype(np.float32)[:, None]
    x0, y0, z0 = np.float32(1) - x1, np.float32(1) - y1, np.float32(1) - z1
    d = data
    out = ((d[i, zi, yi, xi] * x0 + d[i, zi, yi, xi + 1] * x1) * y0 + (d[i, zi, yi + 1, xi] * x0 + d[i, zi, yi + 1, xi + 1] * x1) * y1) * z0 + \
          ((d[i, zi + 1, yi, xi] * x0 + d[i, zi + 1, yi, xi + 1] * x1) * y0 + (d[i, zi + 1, yi + 1, xi] * x0 + d[i, zi + 1, yi + 1, xi + 1] * x1) * y1) * z1
    out = out.astype(np.float32)
    black = np.all(flat == 0.0, axis=1)                       # srgb.cpp:34-37
    white = np.all(flat == 1.0, axis=1)
    out[black] = (0.0, 0.0, -np.inf)
    out[white] = (0.0, 0.0, np.inf)
    return out.reshape(rgb_in.shape)


def model_eval(coeff, wavelengths):
    """``srgb_model_eval`` (include/mitsuba/render/srgb.h:8-24) on the host, float64 -- for tests and previews."""
    coeff = np.asarray(coeff, dtype=np.float64)
    lam = np.asarray(wavelengths, dtype=np.float64)
    v = (coeff[..., 0:1] * lam + coeff[..., 1:2]) * lam + coeff[..., 2:3]
    with np.errstate(invalid="ignore"):
        r = np.maximum(0.0, 0.5 * v / np.sqrt(v * v + 1.0) + 0.5)
    inf = np.isinf(coeff[..., 2:3])
    return np.where(inf, np.sign(coeff[..., 2:3]) * 0.5 + 0.5, r)


def emitter_coeff_scale(rgb, model=None):
    """RGB emission -> (coefficients, scale): srgb_d65.cpp:36-52 / envmap.cpp:98-110 (normalise so the largest
    component is 0.5, fetch, keep ``2 * hmax`` as the scale that multiplies the D65 spectrum)."""
    rgb = np.asarray(rgb, dtype=np.float32)
    scale = rgb.max(axis=-1) * np.float32(2)
    norm = rgb / np.maximum(np.float32(1e-8), scale)[..., None]
    return fetch(norm, model=model), np.asarray(scale, dtype=np.float32)
