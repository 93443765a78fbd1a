"""profiles/sass/: per-kernel SASS listings of librtw.so (cuobjdump -sass), encodings stripped.
Usage: python tools/sass_split.py   (after `make -C raytracing-weekend_b200`)"""
import os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "raytracing-weekend_b200", "librtw.so")],
                     capture_output=True, text=True, check=True).stdout
out = os.path.join(ROOT, "profiles", "sass")
os.makedirs(out, exist_ok=True)
keep = {"k_traverse_small", "k_shadow_small", "k_traverse_flat", "k_shadow_flat", "k_traverse", "k_shadow", "k_shade_all", "k_regen", "k_stream_fill_bvh", "k_stream_shade_bvh", "k_stream_advance", "k_tail_small", "k_tail_flush", "k_karras",
        "k_refit", "k_sort_scatter", "k_sort_scan", "k_sort_tile_hist", "k_morton", "k_prim_aabb", "k_widen", "k_tonemap", "k_trace", "k_ray_bin", "k_bin_scan",
        "k_ray_place", "k_atrous"}
index = []
for p in re.split(r"\n\s*Function : ", txt)[1:]:
    name = p.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    m = re.search(r"(k_[a-z_]+)(?:<([^>]*)>)?\(", dem)
    short = m.group(1) if m else name[:40]
    if m and m.group(2):   # template instance: k_shade_all<3> -> k_shade_all_3, k_sort_scatter<unsigned long long> -> k_sort_scatter_u64
        arg = m.group(2).replace("unsigned long long", "u64").replace("unsigned int", "u32").replace("(bool)", "").replace("(int)", "")
        short_inst = short + "_" + re.sub(r"[^0-9a-z]+", "_", arg.lower()).strip("_")
    else:
        short_inst = short
    body = re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/", "", p)
    body = "\n".join(l.rstrip() for l in body.split("\n") if l.strip())
    index.append((short_inst, len(re.findall(r"/\*[0-9a-f]{4,5}\*/", body))))
    if short in keep and (short != "k_shade_all" or short_inst == "k_shade_all_3") and not short_inst.endswith("_1") and short_inst != "k_sort_scatter_u32":
        open(os.path.join(out, short + ".sass"), "w").write("Function : " + body + "\n")
with open(os.path.join(out, "INDEX.md"), "w") as f:
    f.write("# SASS listings (cuobjdump -sass librtw.so, sm_100a, nvcc 12.9, -O3 -lineinfo -fmad=false)\n\n"
            "Instruction counts per kernel; full text for the hot kernels beside this file (encodings stripped).\n"
            "No tcgen05 / UTMA instructions are expected: the path has no dense contraction (BASELINE.json north_star).\n\n"
            "| kernel | SASS instructions |\n|---|---|\n" + "\n".join("| %s | %d |" % x for x in index) + "\n\n"
            "Mnemonics the design relies on (grep the listings): `LDG.E.ENL2.256` / `STG.E.ENL2.256` (one 256-bit access per\n"
            "32-byte half of a slot record: k_shade_all, k_regen, k_traverse_small), `ATOMG.E.ADD.F32x4.FTZ.RN` (accumulator flush,\n"
            "k_regen; `.FTZ`: DESIGN.md 3), `MATCH.ANY` (one atomic round trip per warp for all queue classes), `LDCU` / `BRA.U` / `BRXU` (the\n"
            "leaf loop of k_traverse_small / k_shadow_small runs on the uniform datapath: leaf records come from the\n"
            "kernel-parameter bank), `CCTL.E.PF2` (`prefetch.global.L2` of the next chunk's records, k_shade_all).\n")
print(index)
