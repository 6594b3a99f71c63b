"""Generates tests/golden/reference_pins.npz from the reference's OWN executable artefacts for this path.

Runs only where /root/reference exists (this container); the fixture it writes is committed so that the pins hold on
the GPU box too.  Two artefacts (the reference has no tests, no golden vectors and no other runnable code on the path):

  * /root/reference/pyscript/debug_regression.py:28-52  regression(points): the 3x3 normal-equation plane fit
    (a, b, c with a*x + b*y + c*z = 1) that shaders/advanced_ssr/regression.comp:41-72 implements; sse() :54-59 is the
    residual of :76-101.  Imported and called as it is.
  * /root/reference/pyscript/integrate.py:7  the sympy expression D(x; theta, phi_n, a) of the GGX-direction integrand
    behind shaders/advanced_ssr/preintegrate.comp:45-84.  exec'd as it is, lambdified, sampled.

The point sets fed to regression() are what one 8x8 tile of the regression pass sees: a 16x16 frame whose half-res depth
level (the level the pass binds, advanced_ssr.cpp:524) holds D24 codes of a tilted plane (+ seeded noise), reconstructed
in float64 (gbuffer_encode.glsl:53-69, uv = pixel / size as regression.comp:25 has it).

    python tools/make_reference_pins.py
"""
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/pyscript"
FOVY, ZN, ZF = float(np.float32(np.pi / 3)), float(np.float32(0.05)), float(np.float32(80.0))
D24 = 16777215.0


def load_regression_module():
    spec = importlib.util.spec_from_file_location("ref_debug_regression", os.path.join(REF, "debug_regression.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)            # its top-level line only builds a 5x5 sample point set
    return mod


def tile_codes(plane, noise, seed):
    """D24 codes (8x8, the half-res level of a 16x16 frame, aspect 1) of the view-space plane a*x + b*y + c*z = 1."""
    tg = np.tan(FOVY / 2)
    px, py = np.meshgrid(np.arange(8), np.arange(8))
    u, v = px / 8.0, py / 8.0                                   # regression.comp:25, no half-texel offset
    kx, ky = -(2 * u - 1) * tg, -(2 * v - 1) * tg               # view_vec = z * (kx, ky, 1)
    z = 1.0 / (plane[0] * kx + plane[1] * ky + plane[2])
    assert (z < -ZN).all() and (z > -ZF).all(), "plane must lie in front of the camera over the whole tile"
    z = z + np.random.default_rng(seed).normal(0.0, noise, z.shape) if noise else z
    d = ZF / (ZF - ZN) + ZF * ZN / (z * (ZF - ZN))              # gbuffer_encode.glsl:71-73
    return np.round(d * D24).astype(np.uint32)


def tile_points(codes):
    """float64 reconstruct_view_vec of the texels the pass fetches (camera_to_world = identity)."""
    tg = np.tan(FOVY / 2)
    px, py = np.meshgrid(np.arange(8), np.arange(8))
    u, v = px / 8.0, py / 8.0
    d = codes.astype(np.float64) / D24
    z = ZN * ZF / (d * (ZF - ZN) - ZF)
    return np.stack([-(2 * u - 1) * z * tg, -(2 * v - 1) * z * tg, z], axis=-1).reshape(64, 3)


def main():
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (run in the build container)")
    ref = load_regression_module()
    cases = [((0.10, 0.20, -0.50), 0.0, 1), ((-0.30, 0.05, -0.25), 0.0, 2), ((0.0, 0.0, -1.0 / 3.0), 0.0, 3),
             ((0.15, -0.25, -0.40), 0.02, 4), ((-0.05, 0.20, -0.30), 0.05, 5), ((0.40, 0.10, -0.60), 0.01, 6)]
    codes, planes, mse = [], [], []
    for plane, noise, seed in cases:
        k = tile_codes(np.array(plane), noise, seed)
        pts = tile_points(k)
        fit = ref.regression(pts)                               # <- the reference's function
        codes.append(k)
        planes.append(np.asarray(fit, dtype=np.float64))
        mse.append(float(ref.sse(pts, fit)) / 64.0)             # <- the reference's residual, per element as regression.comp:99
    # the reference's own sample (debug_regression.py:61 builds gen_points(5, 1, 1, 1)): its fit must be the plane it was built from
    own = ref.regression(ref.points)
    assert np.allclose(own, [1, 1, 1], atol=1e-9), own

    # ---- integrate.py: exec the file as it is, lambdify D ----
    import sympy
    ns = {}
    exec(compile(open(os.path.join(REF, "integrate.py")).read(), "integrate.py", "exec"), ns)
    D = sympy.lambdify((ns["x"], ns["theta"], ns["phi_n"], ns["a"]), ns["D"], "numpy")
    rng = np.random.default_rng(7)
    n = 4096
    xs = rng.uniform(-np.pi / 2 + 1e-3, np.pi / 2 - 1e-3, n)
    th = rng.uniform(0.0, np.pi, n)
    ph = rng.uniform(0.0, np.pi / 2, n)
    al = rng.uniform(0.05, 0.9, n)                              # brdf.glsl:111 clamps alpha to [0, 0.9]
    dv = D(xs, th, ph, al)

    # ---- LUT texels through the reference's D --------------------------------------------------------------------------
    # With t = tan(x/2), c = sqrt(1 - alpha^2), a = c*cos(theta)*cos(phi_n), b = c*sin(phi_n), L = (b-a)*t + (b+a):
    #     D(x) = cos(x) * (1+t^2)^2 / (1 + t^2 - L^2/2)^2          (half-angle substitution; checked in the test on d_*)
    # so the shader's integrand G2 = (1-t)*L / (1 + t^2 - L^2/2)^2 (preintegrate.comp:77-84, L > 0) equals
    #     (1-t) * L * D(2*atan(t)) / (cos(x) * (1+t^2)^2),
    # and (a, b) is exactly the lookup coordinate of brdf.glsl:121-124.  The LUT texel (X, Y) of an S x S table holds
    # int_{-1}^{1} G2 dt at a = 2(X+.5)/S - 1, b = (Y+.5)/S (preintegrate.comp:53-65): integrate that in float64, 2e5 midpoints,
    # with the denominator taken from the reference's expression.  Only texels inside a^2 + b^2 < c^2 are ever looked up.
    S, alpha = 64, 0.05
    c = np.sqrt(1 - alpha * alpha)
    texels, integrals = [], []
    for X in range(2, S, 7):
        for Y in range(1, S, 9):
            a, b = 2 * (X + 0.5) / S - 1, (Y + 0.5) / S
            if a * a + b * b >= 0.97 * c * c:
                continue
            phi = np.arcsin(b / c)
            theta = np.arccos(np.clip(a / (c * np.cos(phi)), -1, 1))
            n = 200000
            t = -1 + (np.arange(n) + 0.5) * (2.0 / n)
            x = 2 * np.arctan(t)
            L = (b - a) * t + (b + a)
            g = np.where(L > 0, (1 - t) * L * D(x, theta, phi, alpha) / (np.cos(x) * (1 + t * t) ** 2), 0.0)
            texels.append((X, Y))
            integrals.append(g.sum() * (2.0 / n))
    print(len(texels), "LUT texels integrated through the reference's D; range", min(integrals), max(integrals))

    out = os.path.join(ROOT, "tests", "golden", "reference_pins.npz")
    np.savez_compressed(out, reg_codes=np.array(codes), reg_planes=np.array(planes), reg_mse=np.array(mse),
                        reg_cases=np.array([list(p) + [s, q] for p, s, q in cases]),
                        d_x=xs, d_theta=th, d_phi=ph, d_alpha=al, d_value=dv,
                        lut_size=S, lut_alpha=alpha, lut_texels=np.array(texels), lut_integrals=np.array(integrals))
    print("wrote", out, os.path.getsize(out), "bytes")
    for (p, s, q), f, m in zip(cases, planes, mse):
        print("plane", p, "noise", s, "->", np.round(f, 6), "mse", m)


if __name__ == "__main__":
    main()
