#!/usr/bin/env python
"""Write SASS listings of the collective kernels + an opcode summary to profiles/sass/ (runs without a GPU)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "torch_distlearn_b200", "lib", "libdistlearn_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
WANT = {
    "sgd_twoshot_kernelILi8ELb1ELb0E": "fused_sgd_nvls_w8", "sgd_twoshot_kernelILi8ELb0ELb0E": "fused_sgd_twoshot_w8",
    "sgd_oneshot_kernelILi8ELb0E": "fused_sgd_oneshot_w8", "fused_ea_kernelILi8ELb1ELb0E": "fused_ea_nvls_w8",
    "fused_ea_kernelILi8ELb0ELb0E": "fused_ea_twoshot_w8", "allreduce_twoshot_kernelILi8ELb1ELb0E": "allreduce_nvls_w8",
    "async_ea_client_kernel": "async_ea_client", "broadcast_kernelILi2ELb0E": "broadcast_nvls",
    "sgd_twoshot_kernelILi1ELb0ELb0E": "fused_sgd_w1",
    "tc_gemm_ts_kernelILi128E": "tc_gemm_ts_bn128", "tc_gemm_kernelILi128E": "tc_gemm_smemA_bn128",
    "bn_relu_pool_bwd_apply_kernel": "bn_relu_pool_bwd_apply", "bn_stats_kernel": "bn_stats",
    # round 2
    "9ll_kernelILi1ELi0ELb0E": "ll_fused_sgd", "9ll_kernelILi0ELi0ELb0E": "ll_allreduce_f32", "9ll_kernelILi3ELi0ELb0E": "ll_broadcast",
    "sgd_pipe_kernelILi8ELb0E": "fused_sgd_hybrid_w8", "sgd_light_kernelILi8ELb0E": "fused_sgd_light_w8",
    "allreduce_pipe_kernelILi8ELb0E": "allreduce_hybrid_w8", "fused_ea_pipe_kernelILi8ELb0E": "fused_ea_hybrid_w8",
    "broadcast_wide_kernelILb1ELb0E": "broadcast_wide_mc", "broadcast_wide_kernelILb0ELb0E": "broadcast_wide_p2p",
    "mnist_step_kernelILb1E": "mnist_step_form2", "tc_gemm_sk_kernelILi128E": "tc_gemm_streamk_bn128",
    "litmus_kernelILb0E": "litmus",
}
KEYS = ("UCGABAR", "MAPA", "LD.E", "ST.E", "LDL", "STL", "LDGMC", "REDG", "RED.", "STG", "LDG", "ATOM", "MEMBAR", "CCTL", "ERRBAR", "UTC", "LDTM", "STTM", "UTMA", "UBLKCP", "SYNCS")


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", txt)[1:]
    summary = []
    for f in funcs:
        name = f.split("\n", 1)[0].strip()
        for k, out in WANT.items():
            if k in name:
                dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
                ops = collections.Counter(m.group(1) for m in re.finditer(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Za-z0-9_.]+)", f))
                key = {o: c for o, c in ops.items() if any(t in o for t in KEYS)}
                open(os.path.join(OUT, out + ".sass"), "w").write("Function : " + f)
                summary.append((out, dem, key))
    with open(os.path.join(OUT, "README.md"), "w") as fh:
        fh.write("# SASS listings of the collective and tensor-core kernels (`cuobjdump -sass libdistlearn_b200.so`, sm_100a)\n\n"
                 "Regenerate with `python scripts/dump_sass.py`.  Memory instructions per kernel (opcode: static count).\n"
                 "`LDGMC.E.ADD.F32x4...` = `multimem.ld_reduce` (in-switch reduction); `STG.E.128.STRONG.SYS` on the multicast\n"
                 "mapping = `multimem.st`; `REDG.E.ADD.F32x4...SYS` = vector atomics over NVLink; `LDG.E.NA.128` / `STG.E.NA.128`\n"
                 "(`L1::no_allocate`) on peer-mapped pointers = P2P loads/stores; `LDG/STG.E.64.STRONG.SYS` + `MEMBAR.ALL.SYS` +\n"
                 "`CCTL.IVALL` = the release/acquire flags of the signal pad; `UTCHMMA` = `tcgen05.mma`, `UTMALDG` = TMA tensor loads, `LDTM`/`STTM` = `tcgen05.ld`/`tcgen05.st` (tensor memory),\n`UTCBAR` = `tcgen05.commit`, `SYNCS.*` = mbarrier operations; `LD.E.128`/`ST.E.128` = generic-address accesses to the local\nheap (parameters, momentum); `LDL/STL` = the stack copy of `dl_comm_t.base[]` that dynamic rank indexing needs.\n\n")
        for out, dem, key in sorted(summary):
            fh.write(f"## {out}.sass\n`{dem}`\n\n")
            for o, c in sorted(key.items()):
                fh.write(f"- `{o}`: {c}\n")
            fh.write("\n")


if __name__ == "__main__":
    main()
