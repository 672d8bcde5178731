#!/usr/bin/env python3
"""Per-kernel SASS listings and a mnemonic table of lib/libb200cnn_cuda.so (no GPU needed).

    python scripts/sass_listing.py            -> profiles/sass/<kernel>.sass (hot kernels, full listing)
                                                 profiles/sass/kernel_mnemonics.txt (every kernel, counts)

What to look for (B200_PROFILING.md): UTC*MMA = tcgen05.mma (".2CTA" = cta_group::2), LDTM = tcgen05.ld, UTMALDG = TMA
tensor loads, UTCBAR = tcgen05.commit (".MULTICAST" = both CTAs' barriers), LDGSTS = cp.async, DMMA = FP64 tensor
cores, RED/ATOM ... .SYS = system-scope reductions on peer memory, ST/LD on peer pointers are ordinary STG/LDG.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "lib", "libb200cnn_cuda.so")
OUT = os.path.join(ROOT, "profiles", "sass")
HOT = ["tc_block_cs_kernel", "tc_block_pair_kernel", "tc_block_db_kernel", "tc_block_swapped_kernel",
       "tc_gemm_kernel", "dgemm_rows_pipelined_kernel", "flag_begin_kernel",
       "flag_barrier_kernel", "flag_wait_kernel", "flag_signal_kernel", "peer_reduce", "host_flag_store_kernel"]
SKIP = ["tc_gemm_kernelILb0ELi128", "tc_gemm_kernelILb1ELi128"]     # 128-wide twins of the listed 256-wide kernels
KEEP = re.compile(r"^(UTC\w*MMA|UTCBAR|UTMALDG|UTMASTG|UBLKCP|LDTM|STTM|LDGSTS|SYNCS|DMMA|HMMA|RED|ATOM|MEMBAR|ERRBAR|CCTL|"
                  r"LDG|STG|LDS|STS|BAR|FFMA|MUFU|UCGABAR|ST|LD)(\.|$)")


def demangle(names):
    p = subprocess.run(["cu++filt"] + names, capture_output=True, text=True)
    return p.stdout.strip().split("\n") if p.returncode == 0 else names


def main():
    if not os.path.exists(LIB):
        sys.exit(f"{LIB} not built (python -m b200cnn.ops._build)")
    os.makedirs(OUT, exist_ok=True)
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
        elif cur is not None:
            kernels[cur].append(line)
    names = list(kernels)
    pretty = dict(zip(names, demangle(names)))
    table = ["cuobjdump -sass lib/libb200cnn_cuda.so — mnemonic counts per kernel "
             "(tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG, tcgen05.commit -> UTCBAR, cp.async -> LDGSTS, "
             "mbarrier -> SYNCS, DMMA = FP64 tensor).  Full listings of the hot kernels: profiles/sass/*.sass", ""]
    written = []
    for k, lines in kernels.items():
        cnt = collections.Counter()
        for l in lines:
            m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", l)
            if m and KEEP.match(m.group(1)):
                cnt[m.group(1)] += 1
        table.append("== " + pretty[k])
        table.append("   " + ", ".join(f"{a}:{b}" for a, b in sorted(cnt.items())))
        short = next((h for h in HOT if h in k), None)
        if short and not any(x in k for x in SKIP):
            suffix = ""
            t = re.search(r"I(L[bi]\d+E)+", k)            # template arguments, e.g. ILb1EE / ILb0ELi256EE
            if t:
                suffix = "_" + "_".join(re.findall(r"L[bi](\d+)E", t.group(0)))
            fn = os.path.join(OUT, f"{short}{suffix}.sass")
            with open(fn, "w") as f:
                f.write(f"// {pretty[k]}\n// cuobjdump -sass lib/libb200cnn_cuda.so (sm_100a)\n")
                # instruction lines only, without the encoding words (the second line of each instruction and the
                # trailing /* 0x... */): address, predicate, mnemonic, operands
                body = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip()) for l in lines
                        if re.search(r"/\*[0-9a-f]{4}\*/", l) or l.strip().startswith(".")]
                f.write("\n".join(body) + "\n")
            written.append(os.path.basename(fn))
    with open(os.path.join(OUT, "kernel_mnemonics.txt"), "w") as f:
        f.write("\n".join(table) + "\n")
    print(f"{len(kernels)} kernels -> profiles/sass/kernel_mnemonics.txt; listings: {', '.join(sorted(written))}")


if __name__ == "__main__":
    main()
