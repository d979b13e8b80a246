#!/usr/bin/env python
"""Regenerates docs/sass/summary.txt: per kernel of libdtf_b200_cuda.so, the count of the SASS
mnemonics that prove which hardware path it uses (tcgen05 -> UTCHMMA / LDTM / UTCBAR, TMA ->
UTMALDG, mbarrier -> SYNCS, peer atomics -> REDG / ATOMG ... .SYS). Runs on the CPU box."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributedtensorflow_b200", "lib", "libdtf_b200_cuda.so")
KEY = re.compile(r"^(UTC|UTMA|LDTM|STTM|SYNCS|REDG|ATOMG|ATOMS|MEMBAR|UCGABAR|ACQBULK|ERRBAR|CCTL|HMMA|REDUX|ELECT|UBLKCP)")
GLOBAL = re.compile(r"^(LDG|STG)")


# one full listing per kernel family: the instance the benchmarks actually run
FAMILIES = {
    "gemm_umma_kernel_64_Kmajor_MNmajor": "void dtf::gemm_umma_kernel<64, false, true>",
    "gemm_umma_kernel_128_MNmajor_MNmajor": "void dtf::gemm_umma_kernel<128, true, true>",
    "gemm_umma_pair_kernel_Kmajor_MNmajor": "void dtf::gemm_umma_pair_kernel<false, true>",
    "conv2_fwd_halo_kernel": "dtf::conv2_fwd_halo_kernel",
    "conv2_dgrad_halo_kernel": "dtf::conv2_dgrad_halo_kernel",
    "conv2_fwd_kernel": "dtf::conv2_fwd_kernel",
    "conv2_dgrad_kernel": "dtf::conv2_dgrad_kernel",
    "conv2_wgrad_kernel": "dtf::conv2_wgrad_kernel",
    "conv1_fwd_kernel": "dtf::conv1_fwd_kernel",
    "conv1_bwd_kernel": "dtf::conv1_bwd_kernel",
    "head_fused_kernel_sparse": "void dtf::head_fused_kernel<false>",
    "push_apply_kernel_adam": "void dtf::push_apply_kernel<3>",
    "mailbox_push_kernel": "dtf::mailbox_push_kernel",
    "mailbox_snapshot_kernel": "dtf::mailbox_snapshot_kernel",
    "mailbox_snapshot_sync_kernel": "dtf::mailbox_snapshot_sync_kernel",
    "mailbox_drain_kernel_adam": "void dtf::mailbox_drain_kernel<3>",
    "sync_token_wait_kernel": "dtf::sync_token_wait_kernel",
    "sync_round_kernel_adam": "void dtf::sync_round_kernel<3>",
    "pull_bulk_kernel": "dtf::pull_bulk_kernel",
    "step_report_kernel": "dtf::step_report_kernel",
}


def dump_listings(out: str) -> int:
    """docs/sass/<family>.sass: the complete SASS of one instance per kernel family."""
    blocks, name, cur = {}, None, []
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                blocks[name] = cur
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            cur = [line]
        elif name:
            cur.append(line)
    if name:
        blocks[name] = cur
    n = 0
    for fname, kname in FAMILIES.items():
        if kname not in blocks:
            print("no kernel named", kname)
            continue
        # keep the instruction text, drop the encoding column (halves the size, loses nothing readable)
        body = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", l.rstrip()) for l in blocks[kname]]
        body = [l for l in body if l.strip()]
        open(os.path.join(ROOT, "docs", "sass", fname + ".sass"), "w").write(
            "// %s - cuobjdump -sass libdtf_b200_cuda.so (sm_100a); regenerate with bench/sass_summary.py\n" % kname
            + "\n".join(body) + "\n")
        n += 1
    return n


def main() -> None:
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    print("full listings:", dump_listings(out))
    kernels, name = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            kernels[name] = [0, collections.Counter(), collections.Counter()]
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]*)", line)
        if m and name:
            op = m.group(1)
            kernels[name][0] += 1
            if KEY.match(op):
                kernels[name][1][op] += 1
            elif GLOBAL.match(op):
                kernels[name][2][op] += 1
    lines = ["# SASS evidence per kernel (cuobjdump -sass libdtf_b200_cuda.so, sm_100a); regenerate with bench/sass_summary.py",
             "# tcgen05.mma -> UTCHMMA (.2CTA = cta_group::2), tcgen05.ld -> LDTM, TMA -> UTMALDG, tcgen05.commit -> UTCBAR,",
             "# mbarrier -> SYNCS.*, cluster barrier -> UCGABAR; legacy mma.sync would show as HMMA (none expected)", ""]
    hmma = 0
    for k, (n, key, glob) in kernels.items():
        hmma += sum(v for o, v in key.items() if o.startswith("HMMA"))
        lines.append(f"{k}  [{n} instr]")
        lines.append("    " + ", ".join(f"{o} x{c}" for o, c in key.most_common()))
        lines.append("    global: " + ", ".join(f"{o} x{c}" for o, c in glob.most_common()))
    lines.append("")
    lines.append(f"# kernels: {len(kernels)}; HMMA (mma.sync) instructions: {hmma}")
    path = os.path.join(ROOT, "docs", "sass", "summary.txt")
    open(path, "w").write("\n".join(lines) + "\n")
    print(path, len(kernels), "kernels, HMMA:", hmma)


if __name__ == "__main__":
    main()
