#!/usr/bin/env python
"""SASS of every kernel on the timed path, from the built library, into profiles/sass_<tag>_<kernel>.txt (mnemonics only, no
encodings) plus profiles/sass_<tag>_index.md (instruction count and the opcodes that identify the pipe: IMMA / UTCIMMA /
UTMALDG / LDTM / IDP / LDGSTS ...).  usage: dump_sass.py <tag>   (run here: cuobjdump needs no GPU)"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
lib = os.path.join(ROOT, "visualslam_b200", "libvslam_b200.so")
WANT = [("octave_mma_k<2, vslam::MmaStaticGeom<64, 11, 13, 17, 21, 25, 31>, 64, false>", "octave_mma_o0"),
        ("octave_mma_k<3, vslam::MmaStaticGeom<64, 21, 25, 31, 39, 49, 63>, 64, false>", "octave_mma_o1"),
        ("octave_mma_k<5, vslam::MmaStaticGeom<64, 39, 49, 63, 79, 99, 123>, 64, false>", "octave_mma_o2"),
        ("octave_mma_k<9, vslam::MmaStaticGeom<32, 79, 99, 123, 155, 195, 245>, 32, false>", "octave_mma_o3"),
        ("octave_mma_k<2, vslam::MmaStaticGeom<64, 11, 13, 17, 21, 25, 31>, 64, true>", "octave_mma_o0_mat"),
        ("octave_umma_k<5, 3, 80, 2>", "octave_umma_o0"), ("octave_umma_k<6, 4, 112, 1>", "octave_umma_o1"),
        ("harris_fast_k<2, 3>", "harris_fast_p2"), ("harris_march_k<2>", "harris_march_p2"), ("upsample2x_k(", "upsample2x"),
        ("reflect_pad_k(", "reflect_pad"), ("reflect_pad_rows_k(", "reflect_pad_rows"), ("seg_reduce_k(", "seg_reduce"), ("seg_spine_k(", "seg_spine"),
        ("seg_apply_k(", "seg_apply"), ("seg_gather_k(", "seg_gather"), ("frame_counts_k(", "frame_counts"), ("grad_all_k(", "grad_all"),
        ("grad_k(", "grad"), ("octave_fused_k<256>", "octave_fused_256"), ("dog_detect_k(", "dog_detect"), ("orient_hist_k(", "orient_hist"), ("sift_k(", "sift")]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
funcs = {}
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); funcs[cur] = []; continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?)\s*;?\s*/\* 0x", line)
    if m and cur:
        funcs[cur].append((m.group(1), m.group(2).rstrip(" ;")))
names = {f: subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip() for f in funcs}
KEY = ("IMMA", "UTCIMMA", "UTMALDG", "LDTM", "IDP", "LDGSTS", "LDSM", "SHFL", "VIMNMX", "I2F", "ATOM", "RED", "STG", "LDG", "BAR", "SYNCS")
idx = [f"# SASS of the timed-path kernels ({tag}), from libvslam_b200.so (nvcc 12.9, sm_100a)", "", "| kernel | file | instructions | identifying opcodes |", "|---|---|---|---|"]
for pat, short in WANT:
    hit = [f for f, n in names.items() if pat in n]
    if not hit:
        idx.append(f"| {pat} | (not in this build) | | |"); continue
    f = hit[0]
    ins = funcs[f]
    ops = collections.Counter()
    for _, t in ins:
        w = t.split()
        op = (w[1] if w and w[0].startswith("@") and len(w) > 1 else w[0]) if w else ""
        ops[op.split(".")[0]] += 1
    fn = f"sass_{tag}_{short}.txt"
    with open(os.path.join(ROOT, "profiles", fn), "w") as fh:
        fh.write(f"// {names[f]}\n// {len(ins)} instructions; cuobjdump -sass, encodings stripped\n")
        for a, t in ins:
            fh.write(f"/*{a}*/ {t}\n")
    keyops = ", ".join(f"{k} {ops[k]}" for k in KEY if ops.get(k))
    idx.append(f"| `{names[f][:110]}` | {fn} | {len(ins)} | {keyops} |")
open(os.path.join(ROOT, "profiles", f"sass_{tag}_index.md"), "w").write("\n".join(idx) + "\n")
print("\n".join(idx))
