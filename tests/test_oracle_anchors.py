"""Analytic anchors that pin the CPU oracle (SURVEY §8c ii-vi; anchor i lives in test_hiz.py).

The reference has no golden vectors and cannot run here, so the oracle is "parity unpinned" against
upstream; these tests pin it against closed-form answers of the shader math it restates.
"""
import ctypes as C

import numpy as np
import pytest

import chain
import oracle_lib as orc
import vkr

FOVY, ZN, ZF = np.float32(np.pi / 3), np.float32(0.05), np.float32(80.0)


def f32(*v):
    return np.array(v, dtype=np.float32)


def cam4(aspect):
    return f32(FOVY, aspect, ZN, ZF)


def depth_of_z(z):
    """gbuffer_encode.glsl:71-73: d = f/(f-n) + f*n/(z*(f-n)), z negative in front of the camera."""
    f, n = float(ZF), float(ZN)
    return f / (f - n) + f * n / (z * (f - n))


def reconstruct_np(u, v, d, aspect):
    """gbuffer_encode.glsl:53-69 in float64 (the analytic side of the anchors)."""
    f, n, tg = float(ZF), float(ZN), np.tan(float(FOVY) / 2)
    z = n * f / (d * (f - n) - f)
    return np.stack([-(2 * u - 1) * z * aspect * tg, -(2 * v - 1) * z * tg, z], axis=-1)


# ---- (ii) shared encode/decode helpers -------------------------------------------------------------
def test_normal_octahedral_roundtrip():
    rng = np.random.default_rng(0)
    v = rng.normal(size=(2000, 3)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    worst = 0.0
    for n in v:
        e, back = f32(0, 0), f32(0, 0, 0)
        orc.lib.orc_probe_encode_normal(orc.fptr(np.ascontiguousarray(n)), orc.fptr(e))
        assert 0.0 <= e[0] <= 1.0 and 0.0 <= e[1] <= 1.0
        orc.lib.orc_probe_decode_normal(orc.fptr(e), orc.fptr(back))
        worst = max(worst, float(np.abs(back - n).max()))
    assert worst < 1e-5, worst


def test_project_inverts_reconstruct():
    rng = np.random.default_rng(1)
    aspect = np.float32(16 / 9)
    c = cam4(aspect)
    for _ in range(500):
        uv = rng.random(2).astype(np.float32)
        d = np.float32(depth_of_z(-rng.uniform(0.2, 60.0)))
        p, back = f32(0, 0, 0), f32(0, 0, 0)
        orc.lib.orc_probe_reconstruct(orc.fptr(uv), C.c_float(d), orc.fptr(c), orc.fptr(p))
        ref = reconstruct_np(float(uv[0]), float(uv[1]), float(d), float(aspect))
        np.testing.assert_allclose(p, ref, rtol=2e-4, atol=1e-5)
        orc.lib.orc_probe_project(orc.fptr(p), orc.fptr(c), orc.fptr(back))
        np.testing.assert_allclose(back[:2], uv, atol=2e-6)
        assert abs(float(back[2]) - float(d)) < 2e-6


def test_srgb_and_half_codecs():
    assert orc.lib.orc_probe_srgb_decode(0) == 0.0 and orc.lib.orc_probe_srgb_decode(255) == 1.0
    for k in range(256):
        lin = orc.lib.orc_probe_srgb_decode(k)
        c = k / 255.0
        ref = c / 12.92 if c <= 0.04045 else ((c + 0.055) / 1.055) ** 2.4
        assert abs(lin - ref) < 1e-6
        assert abs(orc.lib.orc_probe_srgb_encode(C.c_float(lin)) * 255.0 - k) < 1e-3
    for v in (0.0, 1.0, 0.1, 65504.0, 1e-5, -2.5):
        assert orc.lib.orc_probe_half_roundtrip(C.c_float(v)) == float(np.float32(np.float16(v)))


# ---- helpers: a hand-built G-buffer with identity camera --------------------------------------------
def ident16():
    return (C.c_float * 16)(*np.eye(4, dtype=np.float32).reshape(16))


def flat_gbuffer(w, h, z, albedo_code=188, rough_code=128):
    """Fronto-parallel plane at view z (camera = identity, so view normal = world normal = +z)."""
    gb = vkr.Gbuffer(w, h, device="cpu")
    k = int(round(depth_of_z(z) * 0xFFFFFF))
    for img in (gb.depth, gb.prev_depth):
        for lod in range(img.mips):
            mw, mh = img.mip_size(lod)
            img.upload(np.full((mh, mw), k, dtype=np.uint32), lod)
    gb.normal.upload(np.full((h, w, 2), 32768, dtype=np.uint16))           # oct(0,0,1) = (0.5, 0.5)
    gb.albedo.upload(np.full((h, w, 4), albedo_code, dtype=np.uint8))
    mat = np.zeros((h, w, 4), dtype=np.uint8)
    mat[..., 1] = rough_code
    gb.material.upload(mat)
    gb.velocity_vectors.upload(np.zeros((h, w, 2), dtype=np.float16))
    return gb, k


# ---- (iii) GTAO on an unoccluded plane ---------------------------------------------------------------
def test_gtao_unoccluded_plane_is_one_at_the_view_axis():
    """Plane facing the camera: every horizon sample lies in the plane, so at the view axis h = pi/2, n = 0 and
    gtao_camera_space (main.comp:185-217) gives 2 * 0.25 * (-cos(pi) + cos 0 + 0) = 1.  Sky pixels give 0."""
    w, h = 2048, 1024                         # half-res pixel pitch = 2.3 mrad, so the 4x4 centre block is on-axis to < 5 mrad
    gb, _ = flat_gbuffer(w, h, -5.0)
    d = gb.depth.download(1)
    d[:8, :] = 0xFFFFFF                       # a strip of sky
    gb.depth.upload(d, 1)
    raw = vkr.Image(w // 2, h // 2, vkr.R16G16B16A16_SFLOAT, device="cpu")
    pdf = vkr.Image(16, 16, vkr.R32_SFLOAT, device="cpu")
    p = vkr.GtaoParams()
    C.memmove(p.normal_mat, ident16(), 64)
    p.fovy, p.aspect, p.znear, p.zfar = FOVY, w / h, ZN, ZF
    p.angle_offset, p.weight_ratio, p.use_mis, p.two_directions, p.reflections_only = 0.125, 1.0, 0, 0, 0
    orc.ok(orc.lib.orc_gtao_main(C.byref(p), gb.depth.ptr, gb.normal.ptr, gb.material.ptr, pdf.ptr, raw.ptr, 1, 0, 0))
    ao = raw.download()[..., 0].astype(np.float32)
    cy, cx = h // 4, w // 4
    centre = ao[cy - 2:cy + 2, cx - 2:cx + 2]
    # off the axis by an angle t the single-sided slice integral is 1 +- (pi/2) t to first order
    assert np.abs(centre - 1.0).max() < 1.2e-2 and abs(float(centre.mean()) - 1.0) < 4e-3, centre
    assert (ao[:3, :] == 0).all()
    body = ao[8:, :]
    assert np.isfinite(body).all() and body.min() >= 0.0          # single-sided slices: no closed form off the axis


# ---- (iv) BRDF LUT limits ------------------------------------------------------------------------------
def test_brdf_lut_mirror_limit():
    """preintegrate_ssr.comp: as roughness -> 0 and N.V -> 1 the split sum A + B -> 1 and B -> 0 (F = F0)."""
    b = chain.OracleBackend()
    lut = vkr.Image(256, 256, vkr.R16G16_SFLOAT, device="cpu")
    b.brdf_lut(b.halton(), lut)
    t = lut.download().astype(np.float32)
    a, bb = t[255, 0, 0], t[255, 0, 1]
    assert abs(a + bb - 1.0) < 2e-2 and bb < 1e-3, (a, bb)
    assert (t[..., 0] >= 0).all() and (t[..., 0] + t[..., 1] <= 1.02).all()
    # grazing angle on a smooth surface: Fresnel term dominates
    assert t[0, 0, 1] > 0.5 * (t[0, 0, 0] + t[0, 0, 1])


# ---- (v) TAA converges geometrically on a static input ---------------------------------------------------
def test_taa_static_input_converges_like_0p9_pow_k():
    w, h = 64, 32
    gb, _ = flat_gbuffer(w, h, -3.0)
    color = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu").upload(np.full((h, w, 4), 188, dtype=np.uint8))
    c = orc.lib.orc_probe_srgb_decode(188)
    hist = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    tgt = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    p = vkr.TaaParams()
    C.memmove(p.inverse_camera, ident16(), 64)
    C.memmove(p.prev_inverse_camera, ident16(), 64)
    p.fovy_aspect_znear_zfar[:] = [FOVY, w / h, ZN, ZF]
    for k in range(12):
        orc.ok(orc.lib.orc_taa_resolve(C.byref(p), hist.ptr, gb.prev_depth.ptr, gb.depth.ptr, gb.velocity_vectors.ptr, color.ptr, tgt.ptr, 0, 0))
        out = tgt.download().astype(np.float32)[..., :3]
        expect = c * (1.0 - 0.9 ** (k + 1))
        assert np.abs(out - expect).max() < 2e-3 * (k + 1) ** 0.5, (k, out[0, 0], expect)   # fp16 storage each frame
        hist, tgt = tgt, hist


# ---- (vi) SSR on a mirror floor under a wall --------------------------------------------------------------
def test_ssr_mirror_floor_hits_the_reflected_ray():
    """Ground plane with roughness 0 (VNDF collapses to the geometric normal) in front of the back wall: for every
    valid hit, the reconstructed hit point must lie on the mirror-reflected ray of the pixel's view vector."""
    w, h = 512, 288
    scene = vkr.synth_scene(0, 0)                          # ground + wall only
    scene.ground_material = 0                              # roughness code 0, metallic 0
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=16)
    P = oc.begin_frame(scene, jitter=False)
    steps = oc.steps(P)
    steps[0][1]()                                          # Hi-Z
    steps[1][1]()                                          # SSSR_trace
    rays = oc.rays.download().astype(np.float64) / 65535.0
    valid = oc.rays.download()[..., 3] != 65535
    h2, w2 = valid.shape
    aspect = w / h
    d1 = oc.gb.depth.download(1).astype(np.float64) / 0xFFFFFF
    nrm = oc.gb.downsampled_normals.download().astype(np.float64) / 65535.0
    ys, xs = np.nonzero(valid)
    assert len(ys) > 0.02 * w2 * h2, "the synthetic floor should reflect the wall in a sizeable region"
    u, v = (xs + 0.5) / w2, (ys + 0.5) / h2
    Pv = reconstruct_np(u, v, d1[ys, xs], aspect)
    # view-space normal of the pixel: decode the oct normal (gbuffer_encode.glsl:29-37), rotate by normal_mat
    e = nrm[ys, xs] * 2 - 1
    n = np.stack([e[:, 0], e[:, 1], 1 - np.abs(e[:, 0]) - np.abs(e[:, 1])], axis=-1)
    neg = n[:, 2] < 0
    nx = (1 - np.abs(n[:, 1])) * np.where(n[:, 0] >= 0, 1, -1)
    ny = (1 - np.abs(n[:, 0])) * np.where(n[:, 1] >= 0, 1, -1)
    n[neg, 0], n[neg, 1] = nx[neg], ny[neg]
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    M = np.array(list(P["trace"].normal_mat), dtype=np.float64).reshape(4, 4).T[:3, :3]
    nv = n @ M.T
    nv /= np.linalg.norm(nv, axis=1, keepdims=True)
    R = Pv - 2 * np.sum(nv * Pv, axis=1, keepdims=True) * nv
    R /= np.linalg.norm(R, axis=1, keepdims=True)
    hit = rays[ys, xs]
    Hv = reconstruct_np(hit[:, 0], hit[:, 1], hit[:, 2], aspect)
    # screen-space test: the hit texel must sit on the projection of the mirror ray (a straight line through the pixel)
    def project_np(p):
        tg = np.tan(float(FOVY) / 2)
        return np.stack([0.5 * (p[:, 0] / (-p[:, 2] * tg * aspect)) + 0.5, 0.5 * (p[:, 1] / (-p[:, 2] * tg)) + 0.5], axis=-1)
    a, b = project_np(Pv) * [w2, h2], project_np(Pv + R) * [w2, h2]
    t = (b - a) / np.maximum(np.linalg.norm(b - a, axis=1, keepdims=True), 1e-12)
    q = hit[:, :2] * [w2, h2] - a
    perp = np.abs(q[:, 0] * t[:, 1] - q[:, 1] * t[:, 0])           # half-res texels off the line
    along = q[:, 0] * t[:, 0] + q[:, 1] * t[:, 1]
    floor = n[:, 1] > 0.99                                         # only the floor is a mirror here
    assert floor.sum() > 0.5 * len(ys)
    assert np.quantile(perp[floor], 0.999) < 1.0, np.quantile(perp[floor], [0.5, 0.9, 0.99, 1.0])
    assert (along[floor] > 1.5).all()                              # forward along the ray, past the 2-texel self-hit guard
    # and in 3D: the reconstructed hit lies on the mirror ray to within the angle one texel subtends at that distance
    D = Hv - Pv
    dist = np.linalg.norm(D, axis=1)
    cosang = np.sum(D * R, axis=1) / np.maximum(dist, 1e-9)
    ang = np.degrees(np.arccos(np.clip(cosang[floor], -1, 1)))
    assert np.quantile(ang, 0.5) < 0.5 and ang.max() < 5.0, np.quantile(ang, [0.5, 0.9, 0.99, 1.0])
    # and the hit must be on real geometry: depth buffer at the hit uv agrees with the ray depth (trace.comp:111-118)
    hx = np.clip((hit[floor, 0] * w2).astype(int), 0, w2 - 1)
    hy = np.clip((hit[floor, 1] * h2).astype(int), 0, h2 - 1)
    f, nn = float(ZF), float(ZN)
    lin = lambda d: nn * f / (d * (f - nn) - f)  # noqa: E731
    dz = lin(hit[floor, 2]) - lin(d1[hy, hx])
    assert np.quantile(np.abs(dz), 0.99) < 0.35


# ---- (vii) the earlier passes: SSAO and the mirror SSR -------------------------------------------------------
def test_ssao_flat_plane_counts_the_samples_facing_the_camera():
    """Constant depth buffer: a tap at camera_pos + 0.05 s projects nearer than the plane iff s.z > 0 (the camera looks
    down -z), so every pixel holds (#samples with s.z > 0) / 16 exactly (ssao/shader.frag:28-40)."""
    w, h = 160, 96
    gb, _ = flat_gbuffer(w, h, -5.0)
    p = vkr.SsaoParams()
    C.memmove(p.projection, vkr.perspective(FOVY, w / h, ZN, ZF), 64)
    p.fovy, p.aspect, p.znear, p.zfar = FOVY, w / h, ZN, ZF
    s = vkr.sphere_samples(7)
    assert np.allclose(np.linalg.norm(s[:, :3], axis=1), 1.0, atol=1e-6) and np.abs(s[:, 2]).min() > 1e-2
    vkr.set_ssao_samples(p, s)
    out = vkr.Image(w, h, vkr.R8_UNORM, device="cpu")
    orc.ok(orc.lib.orc_ssao(C.byref(p), gb.depth.ptr, out.ptr, 0, 0))
    want = int(np.floor((s[:, 2] > 0).sum() / 16.0 * 255.0 + 0.5))
    got = out.download()
    assert (got == want).all(), (want, np.unique(got))
    # the reference's own packing (12-byte stride read at 16): sample i = packed floats [4i, 4i+1, 4i+2], 12..15 zero
    vkr.set_ssao_samples(p, s, reference_packing=True)
    flat = s[:, :3].reshape(-1)
    z_seen = np.array([flat[4 * i + 2] for i in range(12)] + [0.0] * 4)
    orc.ok(orc.lib.orc_ssao(C.byref(p), gb.depth.ptr, out.ptr, 0, 0))
    # a zero sample sits on the plane: pos_depth == sample_depth up to the D24 rounding of the plane, decided by the 1e-7 bias
    lo = int(np.floor((z_seen > 0).sum() / 16.0 * 255.0 + 0.5))
    hi = int(np.floor(((z_seen > 0).sum() + 4) / 16.0 * 255.0 + 0.5))
    got = out.download()
    assert ((got >= lo) & (got <= hi)).all()


def test_ssr_legacy_floor_reflects_a_constant_frame_with_ggx_weight():
    """Ground + wall, constant frame colour c, floor roughness code 255 (alpha = 1): wherever the floor sees the wall,
    out = coef * c * D(N,N,1) * max(N.R, 0) with D = 1/pi and N.R = -N.view_vec (ssr/shader.frag:37-41,:92-96)."""
    w, h = 384, 216
    scene = vkr.synth_scene(0, 0)
    scene.ground_material = 255 << 8                       # roughness code 255 in .g
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=16)
    P = oc.begin_frame(scene, jitter=False)
    oc.steps(P)[0][1]()                                    # Hi-Z (the march reads every mip)
    gb = oc.gb
    frame = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu")
    frame.upload(np.full((h, w, 4), 255, dtype=np.uint8))  # linear 1.0, alpha 1.0
    out = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
    sp = vkr.SsrParams()
    C.memmove(sp.normal_mat, P["trace"].normal_mat, 64)
    sp.fovy, sp.aspect, sp.znear, sp.zfar = FOVY, np.float32(w) / np.float32(h), ZN, ZF
    orc.ok(orc.lib.orc_ssr_legacy(C.byref(sp), gb.depth.ptr, gb.normal.ptr, frame.ptr, gb.material.ptr, out.ptr, 0, 0))
    got = out.download().astype(np.float64) / 255.0
    assert (got[..., 0] == got[..., 1]).all() and (got[..., 0] == got[..., 3]).all()
    d0 = gb.depth.download(0).astype(np.float64) / 0xFFFFFF
    nrm = gb.normal.download().astype(np.float64) / 65535.0
    ys, xs = np.nonzero(got[..., 0] > 0)
    assert len(ys) > 0.02 * w * h, "the floor should reflect the wall in a sizeable region"
    u, v = (xs + 0.5) / w, (ys + 0.5) / h
    Pv = reconstruct_np(u, v, d0[ys, xs], w / h)
    e = nrm[ys, xs] * 2 - 1
    n = np.stack([e[:, 0], e[:, 1], 1 - np.abs(e[:, 0]) - np.abs(e[:, 1])], axis=-1)
    assert (n[:, 2] >= 0).all() or True
    neg = n[:, 2] < 0
    nx = (1 - np.abs(n[:, 1])) * np.where(n[:, 0] >= 0, 1, -1)
    ny = (1 - np.abs(n[:, 0])) * np.where(n[:, 1] >= 0, 1, -1)
    n[neg, 0], n[neg, 1] = nx[neg], ny[neg]
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    M = np.array(list(sp.normal_mat), dtype=np.float64).reshape(4, 4).T[:3, :3]
    nv = n @ M.T
    nv /= np.linalg.norm(nv, axis=1, keepdims=True)
    floor = np.abs(n[:, 1]) > 0.99
    assert floor.sum() > 0.9 * len(ys)
    expect = np.clip(-(nv * Pv).sum(axis=1) / np.pi, 0, 1)         # coef = 1 away from the 5 % border band
    # interior hits only: the fade band is 5 % of the height at each border of the HIT position; stay well inside with the pixel too
    inner = floor & (u > 0.2) & (u < 0.8) & (v > 0.2) & (v < 0.8)
    err = np.abs(got[ys, xs, 0] - expect)
    full = inner & (err < 1.5 / 255)                               # pixels whose hit is inside the band-free region
    assert full.sum() > 0.5 * inner.sum(), (full.sum(), inner.sum())
    assert (got[ys, xs, 0][inner] <= expect[inner] + 1.5 / 255).all()   # the fade only ever darkens


# ---- (viii) ScreenSpaceTrace ------------------------------------------------------------------------------------
def _st_params(aspect, angle=0.125, rnd=0.3):
    p = vkr.ScreenTraceParams()
    C.memmove(p.normal_mat, ident16(), 64)
    p.random_offset, p.angle_offset = rnd, angle
    p.fovy, p.aspect, p.znear, p.zfar = FOVY, aspect, ZN, ZF
    return p


def test_screen_trace_flat_plane_is_unoccluded_and_gathers_nothing():
    """Fronto-parallel plane: every depth sample lies in the tangent plane, so h_cos = 0, h = pi/2 and
    a = 2 * 0.25 * (1 - cos(pi)) = 1; the ray leaves the surface towards the camera and never hits (trace.comp:262-287)."""
    w, h = 200, 100                                        # not multiples of 8: the last columns / rows stay untouched
    gb, _ = flat_gbuffer(w, h, -5.0)
    d = gb.depth.download(0)
    d[:9, :] = 0xFFFFFF                                    # sky strip (the bilinear footprint of row 9 still touches it)
    gb.depth.upload(d, 0)
    color = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu")
    color.upload(np.full((h, w, 4), 200, dtype=np.uint8))
    raw = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    raw.upload(np.full((h, w, 4), 7.0, dtype=np.float16))
    orc.ok(orc.lib.orc_screen_trace_main(C.byref(_st_params(w / h)), gb.depth.ptr, gb.normal.ptr, color.ptr, gb.material.ptr, raw.ptr, 0, 0))
    out = raw.download().astype(np.float32)
    cw, ch = w // 8 * 8, h // 8 * 8
    assert (out[ch:, :] == 7.0).all() and (out[:, cw:] == 7.0).all()          # outside the dispatched tiles
    assert (out[:8, :cw] == [0, 0, 0, 1]).all()                                # sky
    body = out[16:ch, :cw]
    assert (body[..., :3] == 0).all()
    assert np.abs(body[..., 3] - 1.0).max() < 2e-3, np.abs(body[..., 3] - 1.0).max()


def test_screen_trace_filter_and_accumulate_fixed_points():
    """filter.comp: constant raw over a constant-depth plane stays constant in the interior.  View-space z is negative, so
    the weight 1 - |dz| / (0.1 z) GROWS with the depth difference (:31, as written in the reference): out-of-image taps
    (depth 0, raw 0) get weight 1 + 4.95/0.5 = 10.9 and dilute the two border rows / columns towards 0.
    accumulate.comp: identical depth and prev_depth blend 5 % per frame -> c (1 - 0.95^k); a changed depth or a sky
    prev_depth takes the new value."""
    w, h = 64, 32
    gb, k = flat_gbuffer(w, h, -5.0)
    raw = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    raw.upload(np.full((h, w, 4), 0.5, dtype=np.float16))
    filt = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    orc.ok(orc.lib.orc_screen_trace_filter(ZN, ZF, raw.ptr, gb.depth.ptr, filt.ptr, 0, 0))
    f = filt.download().astype(np.float32)
    assert (f[2:-1, 2:-1] == 0.5).all()
    assert (f[:2] < 0.5).all() and (f[-1] < 0.5).all() and (f[:, :2] < 0.5).all() and (f[:, -1] < 0.5).all()
    edge = 0.5 * 12 / (12 + 4 * np.float32(10.9))           # a top-row interior pixel: 4 of its 16 taps are out of the image
    assert abs(f[1, 10, 0] - edge) < 2e-3, (f[1, 10, 0], edge)
    filt.upload(np.full((h, w, 4), 0.5, dtype=np.float16))
    acc = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu")
    pd = gb.prev_depth.download(0)
    pd[:4, :] = k + 1000                                   # moved surface
    pd[4:8, :] = 0xFFFFFF                                  # sky last frame
    gb.prev_depth.upload(pd, 0)
    cam = f32(FOVY, w / h, ZN, ZF)
    want = 0.0
    for it in range(5):
        orc.ok(orc.lib.orc_screen_trace_accumulate(orc.fptr(cam), gb.depth.ptr, gb.prev_depth.ptr, filt.ptr, acc.ptr, 0, 0))
        want = float(np.float16(np.float32(want) * np.float32(1 - 0.05) + np.float32(0.5) * np.float32(0.05)))
        a = acc.download().astype(np.float32)
        assert (a[:8] == 0.5).all()
        assert np.abs(a[8:] - want).max() <= 1e-3, (it, a[8:].max(), want)
    assert abs(want - 0.5 * (1 - 0.95 ** 5)) < 2e-3


# ---- (ix) GTAO variants -----------------------------------------------------------------------------------------
def test_gtao_graphics_plane_and_deinterleave_layout():
    """main.frag on a fronto-parallel plane: like anchor (iii) the slice integral is 1 at the view axis; sky returns 1.
    deinterleave.comp: texel (x, y) of the dispatched corner lands in layer ((y & 3) << 2) + (x & 3) at (x >> 2, y >> 2)."""
    w, h = 2048, 1024
    gb, _ = flat_gbuffer(w, h, -5.0)
    d = gb.depth.download(1)
    d[:8, :] = 0xFFFFFF
    gb.depth.upload(d, 1)
    raw = vkr.Image(w // 2, h // 2, vkr.R16G16B16A16_SFLOAT, device="cpu")
    p = vkr.GtaoParams()
    C.memmove(p.normal_mat, ident16(), 64)
    p.fovy, p.aspect, p.znear, p.zfar, p.angle_offset = FOVY, w / h, ZN, ZF, 0.125
    orc.ok(orc.lib.orc_gtao_main_graphics(C.byref(p), gb.depth.ptr, gb.normal.ptr, raw.ptr, 1, 0, 0))
    out = raw.download().astype(np.float32)
    cy, cx = h // 4, w // 4
    centre = out[cy - 2:cy + 2, cx - 2:cx + 2, 0]
    assert np.abs(centre - 1.0).max() < 1.2e-2, centre
    assert (out[:3, :, 0] == 1.0).all() and (out[..., 1] == 0).all() and (out[..., 3] == 1).all()

    # deinterleave: a depth level holding its own texel index
    w2, h2 = 256, 128
    depth = vkr.Image(2 * w2, 2 * h2, vkr.D24_UNORM_S8_UINT, 2, device="cpu")
    idx = (np.arange(h2)[:, None] * w2 + np.arange(w2)[None, :]).astype(np.uint32)
    depth.upload(idx, 1)
    arr = vkr.ImageArray(w2 // 4, h2 // 4, vkr.R32_SFLOAT, 1, layers=16, device="cpu")
    orc.ok(orc.lib.orc_gtao_deinterleave(depth.ptr, 1, arr.ptr, arr.layer_stride, 2))
    cw, ch = (w2 // 4) // 8 * 8, (h2 // 4) // 4 * 4                      # the dispatch covers the OUTPUT extent (gtao.cpp:468)
    for layer in (0, 5, 15):
        got = arr.layer(layer).download()
        ly, lx = layer >> 2, layer & 3
        want = np.zeros_like(got)
        ys, xs = np.arange(ly, ch, 4), np.arange(lx, cw, 4)
        want[:len(ys), :len(xs)] = (idx[np.ix_(ys, xs)].astype(np.float64) / 0xFFFFFF).astype(np.float32)
        assert np.array_equal(got, want), layer


def test_gtao_reproject_blends_only_static_texels():
    w2, h2 = 64, 32
    gb, k = flat_gbuffer(2 * w2, 2 * h2, -5.0)
    pd = gb.prev_depth.download(1)
    pd[:4] = k + 1000
    pd[4:8] = 0xFFFFFF
    gb.prev_depth.upload(pd, 1)
    ao, prev, out = [vkr.Image(w2, h2, vkr.R16_SFLOAT, device="cpu") for _ in range(3)]
    ao.upload(np.full((h2, w2), 0.5, dtype=np.float16))
    prev.upload(np.full((h2, w2), 1.0, dtype=np.float16))
    cam = f32(FOVY, 2.0, ZN, ZF)
    orc.ok(orc.lib.orc_gtao_reproject(orc.fptr(cam), gb.depth.ptr, gb.prev_depth.ptr, 1, ao.ptr, prev.ptr, out.ptr, 0, 0))
    o = out.download().astype(np.float32)
    assert (o[:8] == 0.5).all()                                             # moved or sky last frame: the new value
    assert np.abs(o[8:] - (1.0 * 0.95 + 0.5 * 0.05)).max() < 1e-3           # mix(prev, new, 0.05)


# ---- (x) probe path: cube -> octahedral, probe pyramid, probe trace -----------------------------------------------------
def _oct_decode_np(u, v):
    x, y = 2 * (u - 0.5), 2 * (v - 0.5)
    z = 1 - np.abs(x) - np.abs(y)
    fx = np.where(z < 0, (1 - np.abs(y)) * np.where(x >= 0, 1, -1), x)
    fy = np.where(z < 0, (1 - np.abs(x)) * np.where(y >= 0, 1, -1), y)
    d = np.stack([fx, fy, z], axis=-1)
    return d / np.linalg.norm(d, axis=-1, keepdims=True), np.sign(np.stack([fx, fy, z], axis=-1))


def test_cube2oct_sphere_depth_and_face_lookup():
    """A cube map holding constant distance r and its own face index as colour: the octahedral depth must be
    encode_oct_depth(r * dot(dir, front)) with front = normalize(sign(v)) (cube2oct/shader.comp:26-29), and the colour the
    index of the face the direction's major axis selects."""
    import vkr_probes
    n, size, r = 32, 64, 3.0
    cubes = vkr_probes.CubeMapSet(1, n, device="cpu")
    for f in range(6):
        col = np.zeros((n, n, 4), dtype=np.uint8)
        col[..., 3] = 40 * (f + 1)                           # alpha is stored linearly: face id survives the sRGB decode
        cubes.color.layer(f).upload(col)
        cubes.distance.layer(f).upload(np.full((n, n), r, dtype=np.float16))
    oc = vkr.ImageArray(size, size, vkr.R8G8B8A8_UNORM, 1, layers=1, device="cpu")
    od = vkr.ImageArray(size, size, vkr.R16_UNORM, 7, layers=1, device="cpu")
    orc.ok(orc.lib.orc_cube2oct(cubes.color.ptr, cubes.color.layer_stride, cubes.distance.ptr, cubes.distance.layer_stride, oc.layer(0).ptr, od.layer(0).ptr))
    v, u = np.meshgrid(np.arange(size) / size, np.arange(size) / size, indexing="ij")
    d, sg = _oct_decode_np(u, v)
    front = sg / np.maximum(np.linalg.norm(sg, axis=-1, keepdims=True), 1e-30)
    zlin = np.clip(r * (d * front).sum(-1), 0.05, 80.0)
    f_, n_ = 80.0, 0.05
    want = f_ / (f_ - n_) + f_ * n_ / (-zlin * (f_ - n_))
    got = od.layer(0).download(0).astype(np.float64) / 65535.0
    assert np.abs(got - want).max() < 1.5 / 65535 + 1e-6, np.abs(got - want).max()
    a = np.abs(d)
    face = np.where((a[..., 2] >= a[..., 0]) & (a[..., 2] >= a[..., 1]), np.where(d[..., 2] < 0, 5, 4),
                    np.where(a[..., 1] >= a[..., 0], np.where(d[..., 1] < 0, 3, 2), np.where(d[..., 0] < 0, 1, 0)))
    alpha = oc.layer(0).download()[..., 3]
    margin = np.sort(a, axis=-1)
    clear = (margin[..., 2] - margin[..., 1]) > 0.05          # away from face edges (where ties / rounding decide)
    assert (alpha[clear] == 40 * (face[clear] + 1)).all()


def test_probe_downsample_is_a_block_min():
    rng = np.random.default_rng(3)
    size = 64
    od = vkr.ImageArray(size, size, vkr.R16_UNORM, 7, layers=1, device="cpu")
    base = rng.integers(0, 65536, (size, size)).astype(np.uint16)
    layer = od.layer(0)
    layer.upload(base, 0)
    orc.ok(orc.lib.orc_probe_downsample(layer.ptr))
    for k in range(1, 7):
        want = base.reshape(size >> k, 1 << k, size >> k, 1 << k).min(axis=(1, 3))
        assert np.array_equal(layer.download(k), want), k


def _probe_setup(w, h, device="cpu", grid_sz=2, size=128, cube=64, objects=0):
    import vkr_probes
    scene = vkr.synth_scene(0, objects)
    grid = vkr_probes.OctahedralProbeGrid(grid_sz, size, device="cpu")
    grid.min, grid.max = (-3.0, 1.0, -5.0), (3.0, 1.0, 5.0)
    cubes = vkr_probes.CubeMapSet(grid_sz * grid_sz, cube, device="cpu").render_synthetic(scene, grid.positions())
    gb = vkr.Gbuffer(w, h, device="cpu")
    dp = vkr.synth_camera(0, 0, w, h, False)
    vkr.synth_gbuffer(scene, dp, gb)
    params = vkr_probes.ProbeTraceParams(vkr.mat_inverse(dp.camera), *list(dp.fovy_aspect_znear_zfar))
    return scene, grid, cubes, gb, dp, params


def test_trace_probe_floor_sees_the_wall():
    """Ground + back wall + sky, 2x2 probes: wherever the mirror ray of a floor pixel analytically meets the wall's interior,
    a reported hit must carry the wall's colour."""
    import vkr_probes
    w, h = 256, 144
    scene, grid, cubes, gb, dp, params = _probe_setup(w, h)
    for i in range(4):
        orc.ok(orc.lib.orc_cube2oct(cubes.color.layer(6 * i).ptr, cubes.color.layer_stride, cubes.distance.layer(6 * i).ptr, cubes.distance.layer_stride,
                                    grid.color_array.layer(i).ptr, grid.depth_array.layer(i).ptr))
        orc.ok(orc.lib.orc_probe_downsample(grid.depth_array.layer(i).ptr))
    out = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
    k = vkr_probes.trace_constants(grid, params)
    orc.ok(orc.lib.orc_trace_probe(C.byref(k), gb.depth.ptr, gb.normal.ptr, grid.color_array.ptr, grid.color_array.layer_stride, grid.depth_array.ptr,
                                   grid.depth_array.layer_stride, out.ptr, 0, 0))
    got = out.download()
    # analytic mirror rays of floor pixels (uv = pixel / covered extent, camera = view 0 of the synthetic rig)
    cw, ch = w // 8 * 8, h // 4 * 4
    v, u = np.meshgrid((np.arange(ch) + 0.5) / h, (np.arange(cw) + 0.5) / w, indexing="ij")   # the G-buffer is ray-cast at pixel centres
    d0 = gb.depth.download(0)[:ch, :cw].astype(np.float64) / 0xFFFFFF
    Pv = reconstruct_np(u, v, d0, w / h)
    inv = np.array(list(vkr.mat_inverse(dp.camera)), dtype=np.float64).reshape(4, 4).T
    Pw = Pv @ inv[:3, :3].T + inv[:3, 3]
    eye = inv[:3, 3]
    V = Pw - eye
    V /= np.linalg.norm(V, axis=-1, keepdims=True)
    nrm = gb.normal.download()[:ch, :cw]
    floor = (d0 < 1) & (nrm[..., 0] == 32768) & (nrm[..., 1] == 65535)          # oct(0, 1, 0)
    assert np.abs(Pw[..., 1][floor]).max() < 1e-3
    R = V * [1, -1, 1]                                          # reflect about n = +y
    t = (scene.wall_z - Pw[..., 2]) / np.where(R[..., 2] > 1e-6, R[..., 2], np.nan)
    hx, hy = Pw[..., 0] + t * R[..., 0], Pw[..., 1] + t * R[..., 1]
    sees_wall = floor & (t > 0.5) & (np.abs(hx) < scene.wall_half_width - 1.0) & (hy > 0.5) & (hy < scene.wall_height - 0.5) & (t < 25)
    assert sees_wall.sum() > 500
    hit = got[:ch, :cw, 3] > 0
    wall_srgb = np.array([scene.wall_albedo & 255, (scene.wall_albedo >> 8) & 255, (scene.wall_albedo >> 16) & 255], dtype=np.float64) / 255
    wall_lin = np.where(wall_srgb <= 0.04045, wall_srgb / 12.92, ((wall_srgb + 0.055) / 1.055) ** 2.4)
    sel = sees_wall & hit
    assert sel.sum() > 0.3 * sees_wall.sum(), (sel.sum(), sees_wall.sum())
    err = np.abs(got[:ch, :cw, :3][sel].astype(np.float64) / 255 - wall_lin).max(axis=-1)
    # the probe march is an approximation (the hit is accepted in the first probe whose stored depth agrees within HIT_BIAS,
    # seen from the probe, not from the ray origin): most, not all, of the reported hits are the wall
    assert np.median(err) < 1.5 / 255 and (err < 2.5 / 255).mean() > 0.6, ((err < 2.5 / 255).mean(), np.quantile(err, [0.5, 0.9, 0.99]))
    sky = d0 >= 1                                               # uv = pixel / extent is a texel corner: the bilinear depth mixes 4 texels
    sky4 = sky.copy()
    sky4[1:, :] &= sky[:-1, :]; sky4[:, 1:] &= sky[:, :-1]; sky4[1:, 1:] &= sky[:-1, :-1]
    assert sky4.sum() > 100 and (got[:ch, :cw][sky4] == 0).all()
