"""Tier 2 of BASELINE.json measured at the benchmark configuration: the relMSE of an N-spp frame
against a converged frame made from DISJOINT samples (SURVEY.md §8d:
relMSE = mean((I - R)^2 / (R^2 + 1e-2)) over linear RGB), through the C-ABI's rtw_render_accum.

The estimator being matched is the reference's (raygen/raygen.cu:28-87): its variance — not the
implementation — decides at which sample count the 1e-3 target of BASELINE.json is met, so the
ladder of sample counts is measured and reported, not assumed."""
import numpy as np


def relmse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def measure(director, torch, Nx, Ny, ladder=(256, 1024, 4096, 16384), converged_spp=65536, max_depth=20):
    """Renders samples [0, max(ladder)) once (cumulatively: the k-spp frame is the first k samples) and
    samples [max(ladder), max(ladder) + converged_spp) as the converged reference.  Returns
    {"converged_spp", "relmse": {spp: value}, "pixel_pass_rate": {spp: share of pixels with
    |I-R| <= 0.05 + 0.10 R}, "spp_for_1e-3": smallest ladder entry below 1e-3 or None,
    "noise_floor": relMSE of two disjoint halves of the reference}."""
    dev = torch.device("cuda", torch.cuda.current_device())
    acc = torch.zeros(Ny, Nx, 4, device=dev)
    frames, done = {}, 0
    for spp in sorted(ladder):
        director.render_accum(acc.data_ptr(), done, 1, spp - done, max_depth)
        done = spp
        frames[spp] = (acc[..., :3] / spp).cpu().numpy()
    ref_a = torch.zeros(Ny, Nx, 4, device=dev)
    ref_b = torch.zeros(Ny, Nx, 4, device=dev)
    half = converged_spp // 2
    director.render_accum(ref_a.data_ptr(), done, 1, half, max_depth)
    director.render_accum(ref_b.data_ptr(), done + half, 1, converged_spp - half, max_depth)
    ra = (ref_a[..., :3] / half).cpu().numpy()
    rb = (ref_b[..., :3] / (converged_spp - half)).cpu().numpy()
    ref = ((ref_a[..., :3] + ref_b[..., :3]) / converged_spp).cpu().numpy()
    out = {"converged_spp": int(converged_spp), "relmse": {}, "pixel_pass_rate": {}, "spp_for_1e-3": None,
           "noise_floor_two_halves_of_reference": relmse(ra, rb)}
    for spp in sorted(ladder):
        e = relmse(frames[spp], ref)
        out["relmse"][str(spp)] = e
        ok = np.abs(frames[spp] - ref).max(axis=2) <= 0.05 + 0.10 * ref.max(axis=2)
        out["pixel_pass_rate"][str(spp)] = float(ok.mean())
        if out["spp_for_1e-3"] is None and e <= 1e-3:
            out["spp_for_1e-3"] = int(spp)
    return out
