#!/usr/bin/env python3
"""Mnemonic census of every kernel in libpdnn_b200.so (`cuobjdump -sass`) -> profiles/sass/SUMMARY.md, plus the full
listings of the hot kernels. Shows that the GEMM-shaped work runs on tcgen05 (UTCHMMA / LDTM / UTCBAR), is fed by TMA
(UTMALDG) and that no legacy mma.sync (HMMA) is present.

    python tools/sass_census.py
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pytorch_distributed_nn_b200", "lib", "libpdnn_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
COLS = ["UTCHMMA", r"UTCHMMA\.2CTA", "LDTM", "UTMALDG", "UTMALDG.*MULTICAST", "UTMASTG", "UTCBAR", "SYNCS", "UCGABAR", "ACQBULK", "PREEXIT", "HMMA", "STG", r"STG\.E\.ENL2\.256", "LDG", "REDG?|ATOMG"]
KEEP = {          # listing file -> regex on the demangled kernel name
    "conv_halo_kernel.sass": r"pdnn::conv_halo_kernel",
    "conv_halo_wgrad_kernel.sass": r"pdnn::conv_halo_wgrad_kernel",
    "conv_stem_kernel.sass": r"pdnn::conv_stem_kernel",
    "gemm_tcgen05_kernel_CONVK_K2D_256.sass": r"gemm_tcgen05_kernel<\(int\)2, \(int\)0, \(int\)256, \(int\)1>",
    "gemm_tcgen05_kernel_K2D_K2D_256_pair_2cta.sass": r"gemm_tcgen05_kernel<\(int\)0, \(int\)0, \(int\)256, \(int\)3>",
    "gemm_tcgen05_kernel_wgrad_G3_192.sass": r"gemm_tcgen05_kernel<\(int\)1, \(int\)4, \(int\)192, \(int\)1>",
    "ps_bucket_reduce_kernel.sass": r"pdnn::ps_bucket_reduce_kernel<\(bool\)1>",
    "ps_bucket_wait_kernel.sass": r"pdnn::ps_bucket_wait_kernel",
    "ps_push_kernel.sass": r"pdnn::ps_push_kernel",
}


def main():
    if not os.path.exists(LIB):
        sys.exit(f"{LIB} missing: run python csrc/build.py")
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.split("\n")
    parts = re.split(r"\n\s*Function : \S+\n", "\n" + txt)[1:]
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for name, body in zip(names, parts):
        ins = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
        short = re.sub(r"\(int\)", "", name)
        short = re.sub(r"^void ", "", re.sub(r"\(.*", "", short))
        counts = [sum(1 for l in ins if re.search(r"\b(" + c + r")" if c.isalpha() else c, l)) for c in COLS]
        rows.append((short, len(ins), counts))
        for fn, pat in KEEP.items():
            if re.search(pat, name):
                with open(os.path.join(OUT, fn), "w") as f:
                    f.write(f"// {name}\n" + "\n".join(l.rstrip() for l in ins) + "\n")
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as f:
        f.write("# SASS mnemonic census of libpdnn_b200.so (cuobjdump -sass, sm_100a)\n\n"
                "`UTCHMMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG` = TMA load (cp.async.bulk.tensor; `MULTICAST` = the cluster "
                "form), `UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier ops, `UCGABAR` = cluster barrier, `ACQBULK` / `PREEXIT` = griddepcontrol.wait / launch_dependents (programmatic dependent launch), `STG.E.ENL2.256` = 32-byte "
                "stores. No `HMMA` (legacy mma.sync) appears anywhere. Regenerate with `python tools/sass_census.py`.\n\n")
        f.write("| kernel | #instr | " + " | ".join(c.replace("|", "/").replace("\\", "") for c in COLS) + " |\n")
        f.write("|---|---|" + "---|" * len(COLS) + "\n")
        for short, n, counts in sorted(rows, key=lambda r: (-r[2][0], r[0])):
            f.write(f"| `{short}` | {n} | " + " | ".join(str(c) for c in counts) + " |\n")
    print(f"{len(rows)} kernels -> {OUT}/SUMMARY.md; HMMA total = {sum(r[2][COLS.index('HMMA')] for r in rows)}")


if __name__ == "__main__":
    main()
