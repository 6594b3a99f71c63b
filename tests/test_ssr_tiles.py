"""Analytic anchors for the oracle's restatement of the SSSR tile path (classification, indirect trace, tile regression;
reference src/advanced_ssr.cpp:216-302,:440-538).  CPU only."""
import ctypes as C

import numpy as np

import chain
import oracle_lib as orc
import vkr
from test_oracle_anchors import FOVY, ZF, ZN, flat_gbuffer, ident16, reconstruct_np


def classify(material, w2, h2, max_rough=1.0, glossy=0.5):
    n = ((w2 + 7) // 8) * ((h2 + 7) // 8)
    rt, gt = np.full(n, -1, np.int32), np.full(n, -1, np.int32)
    ri, gi = np.array([0, 1, 1], np.uint32), np.array([0, 1, 1], np.uint32)
    orc.ok(orc.lib.orc_sssr_classification(w2, h2, max_rough, glossy, material.ptr, rt.ctypes.data, gt.ctypes.data, ri.ctypes.data, gi.ctypes.data))
    assert ri[1] == 1 and ri[2] == 1 and gi[1] == 1 and gi[2] == 1
    return rt[:ri[0]], gt[:gi[0]]


def material_image(w, h, rough_codes):
    mat = np.zeros((h, w, 4), dtype=np.uint8)
    mat[..., 1] = rough_codes
    return vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu").upload(mat)


def test_classification_splits_tiles_at_the_mean_roughness():
    w, h = 128, 64                                   # rays extent 64 x 32 -> 8 x 4 tiles
    w2, h2 = w // 2, h // 2
    # uniform: every tile lands in one list, in ascending tile index, and the other list stays empty
    for code, mirror in ((0, True), (255, False), (128, True), (200, False)):   # sRGB(128) = 0.216 < 0.5 < sRGB(200) = 0.578
        r, g = classify(material_image(w, h, code), w2, h2)
        full, empty = (r, g) if mirror else (g, r)
        assert list(full) == list(range(32)) and len(empty) == 0, (code, r, g)
    # left half smooth, right half rough, boundary on a tile edge (half-res x = 32): the first rough column reads full-res
    # texels 63|64 (uv = pixel / extent puts the bilinear footprint on the texel corner), mean 61.5/64 -> still glossy
    codes = np.zeros((h, w), dtype=np.uint8)
    codes[:, 64:] = 255
    r, g = classify(material_image(w, h, codes), w2, h2)
    assert sorted(r) == [ty * 8 + tx for ty in range(4) for tx in range(4)]
    assert sorted(g) == [ty * 8 + tx for ty in range(4) for tx in range(4, 8)]
    # the threshold follows g_glossy_value, and max_roughness scales the sampled value first (classification.comp:49)
    r, g = classify(material_image(w, h, 200), w2, h2, max_rough=0.5)
    assert len(r) == 32 and len(g) == 0             # 0.5 * 0.578 < 0.5


def test_classification_counts_pixels_outside_the_image_as_rough():
    """Threads of a partial tile that fall outside `rays` contribute roughness 1 * max_roughness (classification.comp:43-49):
    a mirror surface in a tile that is half outside averages exactly 0.5 and 0.5 < 0.5 is false -> glossy."""
    w, h = 40, 32                                    # rays 20 x 16 -> 3 x 2 tiles, the last column of tiles is half outside
    r, g = classify(material_image(w, h, 0), 20, 16)
    assert sorted(r) == [0, 1, 3, 4] and sorted(g) == [2, 5]
    r, g = classify(material_image(w, h, 0), 20, 16, glossy=0.51)
    assert sorted(r) == [0, 1, 2, 3, 4, 5] and len(g) == 0


def regress(gb, c2w16, aspect):
    w2, h2 = gb.depth.mip_size(1)
    planes = vkr.Image(max(1, gb.depth.width // 16), max(1, gb.depth.height // 16), vkr.R32G32B32A32_SFLOAT, device="cpu")
    cam = np.array([FOVY, aspect, ZN, ZF], dtype=np.float32)
    orc.ok(orc.lib.orc_sssr_tile_regression(c2w16, orc.fptr(cam), gb.depth.ptr, planes.ptr))
    return planes.download().astype(np.float64)


def test_tile_regression_recovers_a_fronto_parallel_plane():
    w, h, z = 128, 96, -4.0
    gb, _ = flat_gbuffer(w, h, z)
    p = regress(gb, ident16(), w / h)
    assert p.shape == (h // 16, w // 16, 4)
    # n . x = 1 with every x on the plane z = -4: n = (0, 0, -1/4); the residual is rounding only
    assert np.abs(p[..., 0]).max() < 2e-3 and np.abs(p[..., 1]).max() < 2e-3, (np.abs(p[..., 0]).max(), np.abs(p[..., 1]).max())
    assert np.abs(p[..., 2] - 1.0 / z).max() < 2e-3, p[..., 2]
    assert p[..., 3].max() < 1e-4, p[..., 3].max()


def test_tile_regression_finds_the_ground_plane_of_the_synthetic_scene():
    """Tiles that see only the ground (world y = 0): positions are camera-relative world offsets (regression.comp:33-36), so
    the fitted plane is (0, -1/cam_height, 0) whatever the camera's orientation."""
    w, h = 512, 288
    scene = vkr.synth_scene(0, 0)
    oc = chain.Chain(w, h, chain.OracleBackend(), luts=(None, None))
    P = oc.begin_frame(scene, jitter=False)
    oc.steps(P)[0][1]()                              # Hi-Z (mip 1 is what the pass reads)
    c2w = vkr.mat_transpose(P["trace"].normal_mat)   # advanced_ssr.cpp:518
    planes = regress(oc.gb, c2w, w / h)
    cam_pos = np.array(list(c2w), dtype=np.float64).reshape(4, 4).T[:3, 3]
    assert cam_pos[1] > 0.5
    # ground tiles: every half-res normal of the 8x8 block points up
    dn = oc.gb.downsampled_normals.download()
    up = np.all(dn == dn[-1, dn.shape[1] // 2], axis=-1)
    th, tw = planes.shape[:2]
    ground = np.array([[up[8 * ty:8 * ty + 8, 8 * tx:8 * tx + 8].all() for tx in range(tw)] for ty in range(th)])
    assert ground.sum() > 0.2 * th * tw
    g = planes[ground]
    expect = -1.0 / cam_pos[1]
    # the normal equations are ill-conditioned for a small patch seen from afar (fp32, cond ~ (distance / patch size)^2)
    err = np.abs(g[:, 1] - expect) / abs(expect)
    assert np.median(err) < 0.02, np.quantile(err, [0.5, 0.9, 1.0])
    near = g[:, 3] < np.median(g[:, 3])
    assert np.abs(g[near, 0]).max() < 0.2 * abs(expect) and np.abs(g[near, 2]).max() < 0.2 * abs(expect)
    assert np.median(g[:, 3]) < 1e-3


def test_trace_indirect_mirror_tiles_hit_the_reflected_ray():
    """Mirror floor under a wall: classification sends the floor's tiles to the reflective list, and every valid hit of the
    indirect trace lies on the screen-space projection of the pixel's mirror ray (cf. anchor vi for the per-pixel trace).
    Pixels of tiles that are not listed stay untouched."""
    w, h = 512, 288
    scene = vkr.synth_scene(0, 0)
    scene.ground_material = 0
    oc = chain.Chain(w, h, chain.OracleBackend(), luts=(None, None))
    P = oc.begin_frame(scene, jitter=False)
    oc.steps(P)[0][1]()
    w2, h2 = w // 2, h // 2
    r, g = classify(oc.gb.material, w2, h2)
    assert len(r) > 0 and len(g) > 0 and len(r) + len(g) == (w2 // 8) * (h2 // 8)
    sentinel = np.full((h2, w2, 4), 12345, dtype=np.uint16)
    oc.rays.upload(sentinel)
    hal = orc.halton23_seq(128)
    orc.ok(orc.lib.orc_sssr_trace_indirect(C.byref(P["trace"]), 0, oc.gb.depth.ptr, oc.gb.downsampled_normals.ptr, oc.gb.material.ptr, orc.fptr(hal),
                                           r.ctypes.data, len(r), oc.rays.ptr))
    raw = oc.rays.download()
    listed = np.zeros((h2, w2), dtype=bool)
    tw = w2 // 8
    for t in r:
        listed[8 * (t // tw):8 * (t // tw) + 8, 8 * (t % tw):8 * (t % tw) + 8] = True
    assert (raw[~listed] == 12345).all() and not (raw[listed] == 12345).all(axis=-1).any()
    dn = oc.gb.downsampled_normals.download()
    # only the floor is a mirror; uv = pixel / extent samples the normal on the texel corner, so keep pixels whose four
    # neighbours all carry the floor's normal
    fl = np.all(dn == dn[-1, w2 // 2], axis=-1)
    fl[1:, 1:] = fl[1:, 1:] & fl[:-1, 1:] & fl[1:, :-1] & fl[:-1, :-1]
    valid = listed & (raw[..., 3] != 65535) & fl
    ys, xs = np.nonzero(valid)
    assert len(ys) > 0.02 * w2 * h2
    rays = raw.astype(np.float64) / 65535.0
    aspect = w / h
    u, v = xs / w2, ys / h2                                            # trace_indirect.comp:53 — no half-texel offset
    # the pixel's own depth is what the shader stores in .w for a valid hit
    Pv = reconstruct_np(u, v, rays[ys, xs, 3], aspect)
    M = np.array(list(P["trace"].normal_mat), dtype=np.float64).reshape(4, 4).T[:3, :3]
    nv = M @ np.array([0.0, 1.0, 0.0])
    nv /= np.linalg.norm(nv)
    R = Pv - 2 * (Pv @ nv)[:, None] * nv
    R /= np.linalg.norm(R, axis=1, keepdims=True)

    def project_np(p):
        tg = np.tan(float(FOVY) / 2)
        return np.stack([0.5 * (p[:, 0] / (-p[:, 2] * tg * aspect)) + 0.5, 0.5 * (p[:, 1] / (-p[:, 2] * tg)) + 0.5], axis=-1)
    a, b = project_np(Pv) * [w2, h2], project_np(Pv + R) * [w2, h2]
    t = (b - a) / np.maximum(np.linalg.norm(b - a, axis=1, keepdims=True), 1e-12)
    q = rays[ys, xs, :2] * [w2, h2] - a
    perp = np.abs(q[:, 0] * t[:, 1] - q[:, 1] * t[:, 0])
    along = q[:, 0] * t[:, 0] + q[:, 1] * t[:, 1]
    assert np.quantile(perp, 0.999) < 1.0, np.quantile(perp, [0.5, 0.9, 0.99, 1.0])
    assert (along > 1.5).all()


def test_trace_indirect_glossy_march_stops_one_level_up():
    """reflection_type 1 marches from view level 1 with 25 cells: on the same mirror floor its hits agree with the mirror march
    to within the coarser cell (2 half-res texels), and being a shorter march it validates no more rays than the mirror one
    does over its own length."""
    w, h = 384, 224
    scene = vkr.synth_scene(0, 0)
    scene.ground_material = 0
    oc = chain.Chain(w, h, chain.OracleBackend(), luts=(None, None))
    P = oc.begin_frame(scene, jitter=False)
    oc.steps(P)[0][1]()
    w2, h2 = w // 2, h // 2
    tiles = np.arange((w2 // 8) * (h2 // 8), dtype=np.int32)
    hal = orc.halton23_seq(128)
    out = []
    for rtype in (0, 1):
        oc.rays.upload(np.zeros((h2, w2, 4), dtype=np.uint16))
        orc.ok(orc.lib.orc_sssr_trace_indirect(C.byref(P["trace"]), rtype, oc.gb.depth.ptr, oc.gb.downsampled_normals.ptr, oc.gb.material.ptr,
                                               orc.fptr(hal), tiles.ctypes.data, len(tiles), oc.rays.ptr))
        out.append(oc.rays.download().astype(np.float64) / 65535.0)
    m, gl = out
    both = (m[..., 3] < 1.0) & (gl[..., 3] < 1.0)
    assert both.sum() > 0.01 * w2 * h2
    d = np.abs(m[both][:, :2] - gl[both][:, :2]) * [w2, h2]
    assert np.quantile(d.max(axis=1), 0.99) <= 4.0, np.quantile(d.max(axis=1), [0.5, 0.9, 0.99, 1.0])
