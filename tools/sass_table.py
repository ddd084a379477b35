"""Regenerates profiles/sass/: per-kernel SASS listings of the hot kernels and the instruction-count
table (UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load, LDTM = tcgen05.ld, LDGMC = multimem.ld_reduce).
Runs on the build machine: cuobjdump needs no GPU.

    python tools/sass_table.py
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_training_dl_b200", "lib", "libdtdl_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")

COLUMNS = [("UTCHMMA", r"\bUTCHMMA"), ("UTMALDG", r"\bUTMALDG"), ("LDTM", r"\bLDTM"),
           ("UTC barrier/alloc", r"\bUTCBAR|\bUTCATOMSWS|\bUTCCP"), ("SYNCS", r"\bSYNCS"), ("HMMA", r"\bHMMA"),
           ("LDG", r"\bLDG\b|\bLDG\."), ("LDGMC", r"\bLDGMC"), ("STG", r"\bSTG"), ("RED/ATOM", r"\bREDG?\b|\bREDG?\.|\bATOM"),
           ("SHFL", r"\bSHFL"), ("MEMBAR", r"\bMEMBAR")]
# mangled-name fragment -> listing file (ILi0E = bf16 instantiation, ILi1E = fp32 / TF32; the bandwidth
# kernels are templated on channels per 16-byte vector: ILi8E = bf16, ILi4E = fp32)
LISTINGS = {"conv_gemm_kernelILi0E": "conv_gemm.sass", "conv_gemm_kernelILi1E": "conv_gemm_tf32.sass",
            "conv_gemm_v2_kernelILi0E": "conv_gemm_v2.sass", "conv_gemm_v2_kernelILi1E": "conv_gemm_v2_tf32.sass",
            "conv_wgrad_kernelILi0E": "conv_wgrad.sass", "conv_wgrad_kernelILi1E": "conv_wgrad_tf32.sass",
            "bn_apply_kernelILi8E": "bn_apply.sass", "bn_apply_stats_kernelILi8E": "bn_apply_stats.sass",
            "bn_bwd_apply_kernelILi8E": "bn_bwd_apply.sass", "bn_bwd_fused_kernelILi8E": "bn_bwd_fused.sass",
            "bn_bwd_fused_kernelILi4E": "bn_bwd_fused_fp32.sass",
            "softmax_ce_kernel": "softmax_ce.sass", "18broadcast_kernel": "broadcast.sass",
            "13reduce_kernelILi8E": "reduce_fused.sass", "nvls_reduce_kernelILi4ELi512E": "nvls_reduce.sass",
            "nvls_reduce16_kernel": "nvls_reduce_bf16.sass", "11sync_kernel": "sync_flags.sass",
            "pack_weights_kernel": "pack_weights.sass"}


def main():
    if not os.path.exists(LIB):
        sys.exit("build the library first (./build.sh)")
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"(?=^\s*Function : )", text, flags=re.M)
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for part in parts:
        m = re.match(r"\s*Function : (\S+)", part)
        if not m:
            continue
        name = m.group(1)
        code = [ln for ln in part.splitlines() if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
        counts = [sum(1 for ln in code if re.search(rx, ln)) for _, rx in COLUMNS]
        rows.append((name, len(code), counts))
        for key, fname in LISTINGS.items():
            if key in name:
                with open(os.path.join(OUT, fname), "w") as f:
                    f.write(part)
    rows.sort(key=lambda r: (-r[2][0], -r[2][1], r[0]))
    with open(os.path.join(OUT, "README.md"), "w") as f:
        f.write("# SASS evidence (`cuobjdump -sass` of libdtdl_b200.so, sm_100a; regenerate with `python tools/sass_table.py`)\n\n")
        f.write("Instruction counts per kernel.  `UTCHMMA` = tcgen05.mma (kind::f16 and kind::tf32 both assemble to it; the "
                "`ILi0E` instantiations of the conv kernels are the bf16 ones, `ILi1E` the fp32 / TF32 ones; for the "
                "bandwidth kernels `ILi8E` = bf16 vectors of 8 channels, `ILi4E` = fp32 vectors of 4), `UTMALDG` = TMA tensor load, `LDTM` = "
                "tcgen05.ld, `SYNCS` = mbarrier operations, `LDGMC` = multimem.ld_reduce (in-switch reduction; the "
                "matching multimem.st is an `STG.E.128.STRONG.SYS` to the multicast address).  No `HMMA` (legacy "
                "mma.sync) anywhere.  The collective kernels dereference peer pointers with plain `LDG`/`STG` (peer "
                "memory is mapped into the address space) plus `.SYS`-scope flag loads and stores.  Full listings "
                "of the main kernels are next to this file.\n\n")
        f.write("| kernel | instrs | " + " | ".join(c for c, _ in COLUMNS) + " |\n")
        f.write("|---|---|" + "---|" * len(COLUMNS) + "\n")
        for name, n, counts in rows:
            f.write(f"| `{name}` | {n} | " + " | ".join(str(c) for c in counts) + " |\n")
    print(f"{len(rows)} kernels -> {OUT}/README.md")
    ptxas_table()


def ptxas_table():
    """`-Xptxas -v` of the last build (build/*.log): registers, spills, static shared memory."""
    import glob
    pat = re.compile(r"Compiling entry function '(\S+)' for 'sm_100a'\nptxas info\s+: Function properties for \S+\n"
                     r"\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n"
                     r"ptxas info\s+: Used (\d+) registers, used (\d+) barriers"
                     r"(?:, (\d+) bytes cumulative stack size)?(?:, (\d+) bytes smem)?")
    rows = []
    for log in sorted(glob.glob(os.path.join(ROOT, "build", "*.log"))):
        src = os.path.basename(log).replace(".o.log", "").replace("csrc_", "").replace("_", "/", 1)
        for m in pat.finditer(open(log).read()):
            name, stack, sst, sld, regs, bars, _cum, smem = m.groups()
            rows.append((src, name, int(regs), int(stack), int(sst), int(sld), int(bars), int(smem or 0)))
    if not rows:
        print("no build logs (run ./build.sh first): ptxas table not written")
        return
    with open(os.path.join(OUT, "ptxas.md"), "w") as f:
        f.write("# `-Xptxas -v` per kernel (sm_100a, `./build.sh`; regenerate with `python tools/sass_table.py`)\n\n")
        f.write("Dynamic shared memory (pipeline stages, up to 224 KB) is chosen at launch and not listed here.\n\n")
        f.write("| source | kernel | registers | stack B | spill st B | spill ld B | barriers | static smem B |\n")
        f.write("|---|---|---|---|---|---|---|---|\n")
        for r in sorted(rows, key=lambda r: (r[0], -r[2])):
            f.write(f"| {r[0]} | `{r[1]}` | {r[2]} | {r[3]} | {r[4]} | {r[5]} | {r[6]} | {r[7]} |\n")
    print(f"{len(rows)} kernels -> {OUT}/ptxas.md")


if __name__ == "__main__":
    main()
