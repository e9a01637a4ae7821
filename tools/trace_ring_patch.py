# development: patch tb_resize.cu with %globaltimer trace points (ring kernel), build to build_ab/libT.so, restore the source
# and the library (the patch is a list of exact-text replacements: it has to follow edits of the kernel)
import subprocess, shutil
p='/root/repo/torchbooster_b200/csrc/tb_resize.cu'
orig=open(p).read()
s=orig
def rep(a,b):
    global s
    assert s.count(a)==1,(a,s.count(a))
    s=s.replace(a,b)
rep("template <typename T, int NT>\n__global__ void __launch_bounds__(kRingMaxGroups * (256 + 32), 1) tb_resize_ring_kernel",
"""__device__ unsigned long long tb_trace[256 * 16];
__device__ __forceinline__ unsigned long long tb_gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TR(k) tb_trace[blockIdx.x * 16 + (k)] = tb_gt()
template <typename T, int NT>
__global__ void __launch_bounds__(kRingMaxGroups * (256 + 32), 1) tb_resize_ring_kernel""")
rep("    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + g.bar_off);        // [ng][slots][2]: full, empty\n",
    "    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + g.bar_off);        // [ng][slots][2]: full, empty\n    if (tid == 0) TR(0);\n    bool tr_first = true, tr_first2 = true;\n")
rep("    tb_grid_dependency_wait();\n    if (warp < n_cons) {", "    tb_grid_dependency_wait();\n    if (tid == 0) TR(1);\n    if (warp < n_cons) {")
rep("        tb_named_barrier(15, nthr);\n    }", "        tb_named_barrier(15, nthr);\n        if (tid == 0) TR(2);\n    }")
rep("                    tb_bulk_g2s(sdst, gsrc0 + (int64_t)r * p.pitch, (uint32_t)ch.seg_len, full);\n                }",
    "                    tb_bulk_g2s(sdst, gsrc0 + (int64_t)r * p.pitch, (uint32_t)ch.seg_len, full);\n                }\n                if (gi == 0 && lane == 0 && tr_first) { TR(3); tr_first = false; }")
rep("            }\n        }\n        return;\n    }\n\n    // ---------------------------------------------------------------------- consumers",
    "            }\n        }\n        if (gi == 0 && lane == 0) TR(8);\n        return;\n    }\n\n    // ---------------------------------------------------------------------- consumers")
rep("        uint32_t eo = 0;                             // element offset of output row yy from dbase\n        tb_named_barrier(1 + grp, g.tc);\n",
    "        uint32_t eo = 0;                             // element offset of output row yy from dbase\n        tb_named_barrier(1 + grp, g.tc);\n        if (tid == 32 && tr_first) TR(9);\n")
rep("                    tb_mbar_wait(full, phase);\n", "                    if (tid == 32 && tr_first) TR(10);\n                    tb_mbar_wait(full, phase);\n                    if (tid == 32 && tr_first) { TR(4); tr_first = false; }\n")
rep("            if (active) {\n#pragma unroll\n                for (int ch = 0; ch < C; ++ch) tb_st_cs<T>",
    "            if (tid == 32 && tr_first2) { TR(5); tr_first2 = false; }\n            if (active) {\n#pragma unroll\n                for (int ch = 0; ch < C; ++ch) tb_st_cs<T>")
rep("            while (yy < nr) emit_zero();\n        }\n    }\n}", "            while (yy < nr) emit_zero();\n        }\n    }\n    if (gt == 32) TR(6 + (grp ? 1 : 0));\n}")
rep("    while (tb_ring_next<NT>(p, g, F, F1, c)) {\n        tb_named_barrier(1 + grp, g.tc);             // every thread of the group is done with the previous chunk's tables\n",
    "    while (tb_ring_next<NT>(p, g, F, F1, c)) {\n        if (tid == 32 && tr_first) TR(11);\n        tb_named_barrier(1 + grp, g.tc);\n        if (tid == 32 && tr_first) TR(12);\n")
rep("            s_ytop[gt] = top;\n        }\n", "            s_ytop[gt] = top;\n        }\n        if (tid == 32 && tr_first) TR(13);\n")
rep("        // one tap encoding per warp: plain if any of its columns has a tap of exactly 1.0, else scaled (top plane\n", "        if (tid == 32 && tr_first) TR(14);\n        // one tap encoding per warp: plain if any of its columns has a tap of exactly 1.0, else scaled (top plane\n")
rep('extern "C" int tb_resize_bilinear_u8_nhwc_to_nchw(', '''extern "C" __attribute__((visibility("default"))) int tb_trace_read(unsigned long long* host, int reset) {
    if (reset) { static unsigned long long z[256 * 16]; return (int)cudaMemcpyToSymbol(tb_trace, z, sizeof(z)); }
    return (int)cudaMemcpyFromSymbol(host, tb_trace, sizeof(unsigned long long) * 256 * 16);
}
extern "C" int tb_resize_bilinear_u8_nhwc_to_nchw(''')
open(p,'w').write(s)
try:
    subprocess.check_call(["python","-m","torchbooster_b200.build"],cwd="/root/repo",stdout=subprocess.DEVNULL)
    shutil.copy("/root/repo/torchbooster_b200/lib/libtorchbooster_b200.so","/root/repo/build_ab/libT.so")
finally:
    open(p,'w').write(orig)
    subprocess.check_call(["python","-m","torchbooster_b200.build"],cwd="/root/repo",stdout=subprocess.DEVNULL)
print("ok")
