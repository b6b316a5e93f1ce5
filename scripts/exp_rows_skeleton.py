#!/usr/bin/env python
"""Build-time experiment behind profiles/r02_gae.md "movement skeleton": how fast is the rows
kernel's DATA MOVEMENT alone?  Patches a copy of csrc/gae.cu (never the shipped file) so that
gae_rows_tma_kernel keeps its bulk loads, stage ring, barriers and bulk stores but

    TS_GAE_EXP=1  touches no shared memory and does no arithmetic (results are garbage),
    TS_GAE_EXP=2  additionally reads its strip, converts f32 -> f64 -> f32 and stages the outputs,

and builds exp_libs/lib_e{0,1,2}.so (0 = shipped source).  On the GPU box:

    for e in 1 2 0; do cp exp_libs/lib_e$e.so tianshou_fork_b200/csrc/libts_b200.so; \
        python scripts/gae_sweep.py --variants 4,5 --iters 500; done
"""
import os
import shutil
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "tianshou_fork_b200", "csrc")

ANCHOR = """        mbar_wait(bar0 + 8u * stage, parity);

        const unsigned char* in = smem + (size_t)stage * L::IN_BYTES;
        const bool live = e0 < n;  // n is a multiple of 16: strips are whole or empty
"""

SKELETON = """        mbar_wait(bar0 + 8u * stage, parity);
#if TS_GAE_EXP == 1 || TS_GAE_EXP == 2
        {
#if TS_GAE_EXP == 2
            const unsigned char* in = smem + (size_t)stage * L::IN_BYTES;
            RewT rw[VEC];
            InT vs[VEC], vn[VEC];
            smem_ld8(reinterpret_cast<const RewT*>(in) + e0, rw);
            smem_ld8(reinterpret_cast<const InT*>(in + L::OFF_VS) + e0, vs);
            smem_ld8(reinterpret_cast<const InT*>(in + L::OFF_VN) + e0, vn);
            const uint2 wt = *reinterpret_cast<const uint2*>(in + L::OFF_TERM + e0);
            const uint2 wc = *reinterpret_cast<const uint2*>(in + L::OFF_TRUNC + e0);
            double d[VEC], r_out[VEC];
#pragma unroll
            for (int k = 0; k < VEC; ++k) { d[k] = (double)rw[k] + (double)vs[k]; r_out[k] = (double)vn[k]; }
            if ((wt.x | wt.y | wc.x | wc.y) == 0x12345678u) d[0] = 0.0;
#endif
            if (tid == 0) bulk_wait_read0();
            __syncthreads();
            if (tid == 0 && p_it < ntiles) issue();
#if TS_GAE_EXP == 2
            smem_st8(s_adv + e0, d);
            smem_st8(s_ret + e0, r_out);
#endif
            fence_proxy_async_smem();
            __syncthreads();
            if (tid == 0) {
                const int64_t v = (int64_t)row * val_stride + (int64_t)c * TILE;
                bulk_s2g(adv + v, smem_u32(s_adv), (uint32_t)n * sizeof(OutT));
                bulk_s2g(ret + v, smem_u32(s_ret), (uint32_t)n * sizeof(OutT));
                bulk_commit();
            }
            if (--c < 0) { c = nchunks - 1; row += gridDim.x; }
            if (++stage == stages) { stage = 0; parity ^= 1u; }
            continue;
        }
#endif

        const unsigned char* in = smem + (size_t)stage * L::IN_BYTES;
        const bool live = e0 < n;  // n is a multiple of 16: strips are whole or empty
"""


def main():
    src = open(os.path.join(CSRC, "gae.cu")).read()
    assert src.count(ANCHOR) == 1, "gae_rows_tma_kernel changed: update ANCHOR"
    out = os.path.join(ROOT, "exp_libs")
    os.makedirs(out, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        work = os.path.join(tmp, "tianshou_fork_b200", "csrc")
        shutil.copytree(CSRC, work, ignore=shutil.ignore_patterns("*.o", "*.so", "shim_build", "*.log"))
        shutil.copytree(os.path.join(ROOT, "include"), os.path.join(tmp, "include"))
        open(os.path.join(work, "gae.cu"), "w").write(src.replace(ANCHOR, SKELETON))
        for e in (0, 1, 2):
            subprocess.check_call(["make", "-s", "-C", work, "clean"])
            subprocess.check_call(["make", "-s", "-C", work, "-j8", f"EXTRA=-DTS_GAE_EXP={e}"])
            shutil.copy(os.path.join(work, "libts_b200.so"), os.path.join(out, f"lib_e{e}.so"))
            print("built", os.path.join(out, f"lib_e{e}.so"))


if __name__ == "__main__":
    main()
