"""Parity metrics and tolerances (SURVEY §8d 'Tolerances to state with results').

Every comparison is CUDA output vs the repo's own CPU restatement (oracle/) on identical input bytes -- not vs the
reference's shaders, which cannot run here; the restatement itself is pinned only where tests/test_reference_pins.py and
tests/test_producer_conventions.py say so.

DEVIATION FROM SURVEY §8d, stated: §8d words the fp bounds as max-abs (GTAO <= 2e-3, TAA <= 4e-3).  The predicates below
bound a QUANTILE instead (|diff| <= bound on >= 99.9 % / 99.99 % of the texels) and keep PSNR as the global bound, because
every one of these images sits behind a discontinuous choice (a reprojection accept / reject, a hit / miss, a horizon
sample entering at |v| < 0.3) that flips on a handful of texels under any arithmetic that is not bit-identical, and a
flipped texel differs by the full range.  So that the outliers stay visible, every metric dict also carries the max-abs
(`max_abs` / `max_lsb` / `x_max`, `w_max`) and the absolute NUMBER of texels beyond the bound (`n_gt`, `n_gt4`, `n_gt1lsb`,
`n_flips`, `n_gt1_texel`), and the committed reports (profiles/r02_parity_*.json) list them per image and frame.
"""
import numpy as np


def psnr(a, b, peak=1.0):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    m = np.isfinite(a) & np.isfinite(b)
    mse = float(np.mean((a[m] - b[m]) ** 2)) if m.any() else 0.0
    return float("inf") if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def compare(name, got, ref):
    """Returns a dict of metrics for image `name` (numpy arrays as Image.download() gives them)."""
    if name in ("depth", "dn", "dv"):           # integer / bit-copy outputs
        g, r = got.view(np.uint16) if got.dtype == np.float16 else got, ref.view(np.uint16) if ref.dtype == np.float16 else ref
        return dict(kind="bits", mismatches=int((g != r).sum()), total=int(g.size))
    if name == "rays":                           # RGBA16_UNORM: xyz = hit uv+z, w = valid ? pixel_depth : 1
        gv, rv = got[..., 3] != 65535, ref[..., 3] != 65535
        flips = float((gv != rv).mean())
        both = gv & rv
        duv = np.abs(got[..., :2].astype(np.int64) - ref[..., :2].astype(np.int64)) / 65535.0
        h, w = got.shape[:2]
        texels = np.maximum(duv[..., 0] * w, duv[..., 1] * h)
        # "all rays" includes invalid ones: the stored position is still the march end point
        return dict(kind="rays", flip_frac=flips, n_flips=int((gv != rv).sum()), max_texels_valid=float(texels[both].max()) if both.any() else 0.0,
                    frac_gt1_texel_all=float((texels > 1.0).mean()), n_gt1_texel=int((texels > 1.0).sum()), max_texels_all=float(texels.max()),
                    valid_frac=float(rv.mean()), total=int(gv.size))
    if got.dtype == np.uint8:                    # unorm8 / srgb8 colour
        d = np.abs(got[..., :3].astype(np.int32) - ref[..., :3].astype(np.int32))
        return dict(kind="u8", frac_le1=float((d <= 1).mean()), n_gt1lsb=int((d > 1).sum()), max_lsb=int(d.max()), total=int(d.size),
                    psnr=psnr(got[..., :3] / 255.0, ref[..., :3] / 255.0))
    g, r = got.astype(np.float32), ref.astype(np.float32)   # fp16 outputs
    if name == "raw_trace":
        # .x = occlusion of the SSR ray, .y = its GGX-direction pdf.  The pdf comes from a LUT whose
        # integrand has poles (reference preintegrate.comp:77-84): values reach 1e4..inf and react to 1 ulp
        # of the lookup coordinate, so .y is judged through the only thing that consumes it, the MIS weight
        # alpha / (alpha*pdf + beta/(2 pi)) of gtao/main.comp:265 (alpha = beta = 0.5).
        w = lambda p: 0.5 / (0.5 * np.nan_to_num(p, nan=0.0, posinf=6.5e4) + 0.5 / (2 * np.pi))  # noqa: E731
        dx = np.abs(g[..., 0] - r[..., 0])
        dw = np.abs(w(g[..., 1]) - w(r[..., 1])) / (2 * np.pi)
        return dict(kind="f16", x_frac_gt=float((dx > 2e-3).mean()), x_n_gt=int((dx > 2e-3).sum()), x_max=float(dx.max()), x_psnr=psnr(g[..., 0], r[..., 0]),
                    w_frac_gt=float((dw > 2e-3).mean()), w_n_gt=int((dw > 2e-3).sum()), w_max=float(dw.max()), total=int(dx.size), nan_mismatch=int((np.isnan(g[..., :2]) != np.isnan(r[..., :2])).sum()))
    if name == "raw":
        g, r = g[..., :2], r[..., :2]
    if name == "taa":
        g, r = g[..., :3], r[..., :3]
    d = np.abs(g - r)
    fin = np.isfinite(d)
    nan_mismatch = int((np.isnan(g) != np.isnan(r)).sum())
    return dict(kind="f16", max_abs=float(d[fin].max()) if fin.any() else 0.0, p999=float(np.quantile(d[fin], 0.999)) if fin.any() else 0.0,
                frac_gt=float((d[fin] > 2e-3).mean()) if fin.any() else 0.0, frac_gt4=float((d[fin] > 4e-3).mean()) if fin.any() else 0.0,
                n_gt=int((d[fin] > 2e-3).sum()), n_gt4=int((d[fin] > 4e-3).sum()), total=int(d.size), psnr=psnr(g, r), nan_mismatch=nan_mismatch)


# name -> predicate on the metrics dict; the numbers are the stated tolerances
TOLERANCES = {
    "depth": lambda m: m["mismatches"] == 0,
    "dn": lambda m: m["mismatches"] == 0,
    "dv": lambda m: m["mismatches"] == 0,
    # hit/miss flag mismatch <= 0.1 % of pixels; hit uv within 1 half-res texel on matching valid pixels
    # (a march whose branch flips on a 1-ulp difference may end elsewhere: allowed on <= 0.01 % of pixels)
    "rays": lambda m: m["flip_frac"] <= 1e-3 and m["frac_gt1_texel_all"] <= 1e-4,
    "raw_trace": lambda m: m["x_frac_gt"] <= 1e-3 and m["w_frac_gt"] <= 1e-3 and m["nan_mismatch"] <= 2,
    # fp16 AO outputs: |diff| <= 2e-3 on >= 99.9 % of texels; PSNR >= 60 dB for filtered / accumulated_ao.
    # raw: PSNR >= 50 dB (not §8d's 60): raw.xy is the UNFILTERED one-slice estimate, where a horizon tap that crosses the
    # thickness break on one side only moves the texel by up to its full range; the 4x4 filter averages that away before
    # anything consumes it, which is why `filtered` keeps the 60 dB bound.  Measured 75-100 dB (profiles/r02_parity_*.json).
    "raw": lambda m: m["frac_gt"] <= 1e-3 and m["psnr"] >= 50.0 and m["nan_mismatch"] <= 2,
    "filtered": lambda m: m["frac_gt"] <= 1e-3 and m["psnr"] >= 60.0,
    "accumulated_ao": lambda m: m["frac_gt"] <= 1e-3 and m["psnr"] >= 60.0,
    # unorm8 reflections: <= 1 LSB on >= 99.9 %, PSNR >= 45 dB
    "reflections": lambda m: m["frac_le1"] >= 0.999 and m["psnr"] >= 45.0,
    "blurred": lambda m: m["frac_le1"] >= 0.999 and m["psnr"] >= 45.0,
    # sRGB8 lighting: <= 1 LSB on >= 99.9 %, PSNR >= 50 dB
    "color": lambda m: m["frac_le1"] >= 0.999 and m["psnr"] >= 50.0,
    # fp16 TAA: |diff| <= 4e-3 on >= 99.99 %, PSNR >= 50 dB
    "taa": lambda m: m["frac_gt4"] <= 1e-4 and m["psnr"] >= 50.0,
    "pdf": lambda m: m["max_rel"] <= 1e-3,
    "brdf": lambda m: m["max_abs"] <= 2e-3,
}


def download_all(img):
    return [img.download(l) for l in range(img.mips)]


def compare_image(name, got_img, ref_img, task=""):
    if name == "raw" and task == "SSSR_trace":
        name = "raw_trace"
    if name == "depth":
        mism = 0
        total = 0
        for lod in range(1, got_img.mips):
            g, r = got_img.download(lod), ref_img.download(lod)
            mism += int((g != r).sum())
            total += g.size
        return dict(kind="bits", mismatches=mism, total=total)
    return compare(name, got_img.download(), ref_img.download())
