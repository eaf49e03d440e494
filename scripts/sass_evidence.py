#!/usr/bin/env python
"""Split `cuobjdump -sass` of libwave3d_b200.so into per-kernel listings of the kernels that matter and
write the mnemonic-count index profiles/sass/README.md (proof of TMA, mbarrier, system-scope flags,
cluster barriers in the shipped binary).  usage: scripts/sass_evidence.py"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mpi_cuda_b200", "_native", "libwave3d_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEEP = {   # file name -> substring of the mangled kernel name
    "pair_step_kernel_64x32x3st_rpt4.sass": "pair_step_kernelINS_7PairCfgILi64ELi4ELi8ELi3ELi1ELi1E",
    "pair_step_kernel_cluster_1x2.sass": "pair_step_kernelINS_7PairCfgILi64ELi4ELi8ELi3ELi1ELi2E",
    "fused_step_kernel_64x32x4st.sass": "fused_step_kernelINS_7TileCfgILi64ELi8ELi4ELi4E",
    "fixup_kernel.sass": "fixup_kernel",
}
INTEREST = re.compile(r"^(UTMALDG|UTMAPF|SYNCS|UCGABAR|CGAERRBAR|MEMBAR|FENCE|ERRBAR|BAR\.|ATOMG|ATOM\.|RED\.|REDG|LDG|STG|LD\.E|ST\.E|LDS|STS|"
                      r"DFMA|DADD|DMUL|ACQBULK|PREEXIT|MAPA|S2UR|LDSM|CCTL)")


def main():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, name = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
        elif name:
            funcs[name].append(line)
    os.makedirs(OUT, exist_ok=True)
    idx = ["# SASS evidence (cuobjdump -sass mpi_cuda_b200/_native/libwave3d_b200.so, sm_100a)\n",
           "Regenerate with `scripts/sass_evidence.py`.  Full listings: " +
           ", ".join(f"`{k}`" for k in KEEP) + ".\n",
           "Mnemonic counts per kernel.  TMA = `UTMALDG.3D`; mbarrier = `SYNCS.*`; generic->async proxy fence after the flag "
           "acquire = `FENCE.VIEW.ASYNC.G`; system-scope flag handshake and P2P stores = `LD/ST/LDG/STG ...STRONG.SYS` + "
           "`MEMBAR.SC.SYS`; thread-block-cluster barrier = `UCGABAR_ARV` / `UCGABAR_WAIT` (cluster variant only).\n"]
    for fn, body in funcs.items():
        for fname, key in KEEP.items():
            if key in fn:
                # keep address + instruction text, drop the encoding words (two thirds of the bytes)
                lines = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", l).rstrip() for l in body]
                lines = [l for l in lines if l.strip()]
                with open(os.path.join(OUT, fname), "w") as fh:
                    fh.write(f"// {fn}\n" + "\n".join(lines) + "\n")
        cnt = collections.Counter()
        n_inst = 0
        for line in body:
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if not m:
                continue
            n_inst += 1
            op = m.group(1)
            if INTEREST.match(op):
                cnt[op] += 1
        if "w3d" not in fn and "error_rows" not in fn:
            continue
        idx.append(f"\n## `{fn}`  ({n_inst} instructions)\n")
        idx.append(", ".join(f"{k} x{v}" for k, v in sorted(cnt.items())))
    with open(os.path.join(OUT, "README.md"), "w") as fh:
        fh.write("\n".join(idx) + "\n")


if __name__ == "__main__":
    main()
