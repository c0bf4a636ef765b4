#!/usr/bin/env python
"""From an `ncu --set full --import-source on` capture of a render kernel to the three text files kept under profiles/:
    <out>_ncu_full_metrics.txt   the metrics DESIGN.md quotes (tools/ncu_metrics.py)
    <out>_by_line.txt            samples / executed instructions / threads per instruction per CUDA source line (tools/ncu_by_line.py)
    <out>_by_phase.txt           the same summed over the phases of the query + integrate loop
usage: ncu_report.py <capture.ncu-rep> <mangled kernel name or unique substring> <out prefix>
The phase ranges are located by anchors in the CURRENT sources (function heads and the `---- T/M/G + E` comments), so the
capture must come from a library built from this tree (the in-tree librt_b200.so is disassembled for the line table)."""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "cpuraytracer_b200", "csrc")


def sh(cmd, **kw):
    return subprocess.run(cmd, shell=True, check=True, capture_output=True, text=True, **kw).stdout


def anchors(path, pats):
    """line numbers (1-based) of the first line matching each regex, in file order"""
    lines = open(path).read().splitlines()
    out = []
    for name, pat in pats:
        hit = next((i + 1 for i, l in enumerate(lines) if re.search(pat, l)), None)
        if hit is None:
            raise SystemExit(f"anchor {name!r} ({pat}) not found in {path}")
        out.append((name, hit))
    out.sort(key=lambda t: t[1])
    return out, len(lines)


def ranges(fname, pats):
    a, n = anchors(os.path.join(CSRC, fname), pats)
    return [f"{name}:{fname}:{lo}-{(a[i + 1][1] - 1) if i + 1 < len(a) else n}" for i, (name, lo) in enumerate(a)]


def main():
    rep, kern, out = sys.argv[1:4]
    tmp = tempfile.mkdtemp(prefix="ncu_report_")
    src_csv, raw_csv, sass = (os.path.join(tmp, f) for f in ("source.csv", "raw.csv", "lib.sass"))
    open(src_csv, "w").write(sh(f"ncu -i {rep} --page source --csv"))
    open(raw_csv, "w").write(sh(f"ncu -i {rep} --page raw --csv"))
    sh(f"cuobjdump -xelf all {os.path.join(CSRC, 'librt_b200.so')}", cwd=tmp)
    cubin = next(f for f in os.listdir(tmp) if f.endswith(".cubin"))
    open(sass, "w").write(sh(f"nvdisasm --print-line-info {cubin}", cwd=tmp))
    names = re.findall(r"^\.text\.(\S+):", open(sass).read(), re.M)
    match = [n for n in names if kern in n]
    if len(match) != 1:
        raise SystemExit(f"kernel {kern!r} matches {match or names}")
    tools = os.path.join(ROOT, "tools")
    open(out + "_ncu_full_metrics.txt", "w").write(sh(f"{sys.executable} {tools}/ncu_metrics.py {raw_csv}"))
    byline = sh(f"{sys.executable} {tools}/ncu_by_line.py {src_csv} {sass} {match[0]} 2000")
    open(out + "_by_line.txt", "w").write(f"# {match[0]}, per CUDA source line (tools/ncu_by_line.py)\n" + "\n".join(byline.splitlines()[:80]) + "\n")
    full = os.path.join(tmp, "byline_all.txt")
    open(full, "w").write("# all lines\n" + byline)
    spec = (ranges("rt_bvh.cuh", [("stage_accel", r"AccelSmem stage_accel"), ("cull_pair2", r"void cull_pair2"), ("slab_consts", r"^struct BoxRay"),
                                  ("box_touched", r"bool box_touched"), ("accel_setup", r"QueryResult fused_query_accel"), ("T", r"// ---- T:"),
                                  ("M", r"// ---- M:"), ("G", r"// ---- G \+ E"), ("E", r"int e = 0; uint32_t cur")])
            + ranges("rt_device.cuh", [("vector_math", r"^struct V3"), ("materials_math", r"XMVector3RefractV"), ("rng_halton", r"Counter-based RNG"),
                                       ("camera_texture", r"^struct DevCamera"), ("sphere_exact", r"bool sphere_exact"), ("ffma2_fadd2", r"u64 pk2\("),
                                       ("brute_cull_lists", r"^struct QueryResult"), ("exact_candidate", r"bool exact_candidate"),
                                       ("brute_cull_pair", r"^struct CullRay "), ("brute_query", r"QueryResult fused_query\("), ("first_hit_multi_query", r"void closest_query_multi")])
            + ranges("rt_kernels.cuh", [("stage_scene", r"SceneView stage_scene"), ("direct_light", r"V3 direct_unshadowed"), ("scatter", r"bool scatter\("),
                                        ("halton_draw", r"void draw_halton"), ("work_items", r"bool item_to_pixel"), ("one_kernel_path_loop", r"void .*rt_render_kernel"),
                                        ("first_hit_kernel", r"^rt_first_hit_kernel\(const RenderParams p\)"), ("path_setup", r"void .*rt_path_kernel"),
                                        ("path_regeneration", r"// ---- regeneration: work items"), ("path_primary_vertex", r"// ---- K1 again for the regenerated"),
                                        ("path_shade_glue", r"// ---- K4 at the vertex in hand"), ("path_query_call", r"fused closest-hit / shadow query ----\s*$"),
                                        ("path_resolve", r"// ---- resolve the pass"), ("path_stats", r"primary rays were counted by stage 1"),
                                        ("tables_kernel", r"Per-call Halton tables on the device"), ("other_kernels", r"^// hdr\[p\] = ")])
            + ["shf_funnel:sm_32_intrinsics:0-99999", "shfl_vote:sm_30_intrinsics:0-99999"])
    phases = sh(f"{sys.executable} {tools}/ncu_by_phase.py {full} " + " ".join(spec))
    open(out + "_by_phase.txt", "w").write(f"# {match[0]}, by phase (tools/ncu_report.py: ranges from anchors in the current sources)\n" + phases
                                           + "# ranges: " + " ".join(spec) + "\n")
    print(phases)


if __name__ == "__main__":
    main()
