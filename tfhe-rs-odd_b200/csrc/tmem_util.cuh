// tmem_util.cuh -- device helpers shared by the tensor-memory blind-rotation kernels: tcgen05.st/ld wrappers (32x32b shape:
// thread l of a warp owns TMEM lane 32*(warp%4)+l), fences, named barriers, mbarriers, the 32x32 split transpose and the
// closed-form rotated digit.
#pragma once
#include "br_core.cuh"

namespace b200tfhe {

#if defined(__CUDA_ARCH__)
DEV void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]),
        "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]),
        "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]),
        "r"(v[31])
        : "memory");
}
DEV void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
DEV void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
DEV void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
DEV void tmem_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
DEV void tmem_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
DEV void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
DEV void named_bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// 32x32 transpose of one warp's register tile through a half-size region (R rows of R/2 + 1 cplx):
// in: x[k1] of lane s = S[k1][s];  out: x[p] of lane l = S[l][brev(p)].
template <int R> DEV void transpose_split(int lane, cplx (&x)[R], cplx *hreg) {
    constexpr int H = R / 2, ROWH = H + 1, LOGR = clog2(R);
    cplx t[H];
    if (lane < H) static_for<0, R>([&](auto k1) { hreg[decltype(k1)::value * ROWH + lane] = x[decltype(k1)::value]; });
    __syncwarp();
    static_for<0, H>([&](auto c) { t[decltype(c)::value] = hreg[lane * ROWH + decltype(c)::value]; });
    __syncwarp();
    if (lane >= H) static_for<0, R>([&](auto k1) { hreg[decltype(k1)::value * ROWH + lane - H] = x[decltype(k1)::value]; });
    __syncwarp();
    static_for<0, R>([&](auto p) {
        constexpr int c = brev(decltype(p)::value, LOGR);
        if constexpr (c < H) x[decltype(p)::value] = t[c];
        else x[decltype(p)::value] = hreg[lane * ROWH + c - H];
    });
    __syncwarp();
}

// (P * X^a)[c] - own for a 64-bit torus and ONE decomposition level, returned as the double value of the balanced digit.
// s8 = 8 * (c - a): bits 3..13 address the coefficient, bit 14 is the sign (X^N = -1).  Closed form of
// BRCore::single_level_digit (exhaustively identical over all 2^32 upper words; the lower word only matters through the
// borrow of the subtraction):  w = (2^31 + 2^sh - 1 - d_hi) >> (sh + 1),  digit = B/2 - w,  sh = 32 - BASE_LOG - 1.
template <int BASE_LOG> DEV double rot_digit_f64(const unsigned char *acc_bytes, uint32_t s8, uint64_t own) {
    constexpr int SH = 32 - BASE_LOG - 1;
    const uint64_t v = *reinterpret_cast<const uint64_t *>(acc_bytes + (s8 & 0x3FF8u));
    const uint32_t neg = (s8 >> 14) & 1u;
    const uint32_t m = 0u - neg;
    const uint64_t d = (v ^ (((uint64_t)m << 32) | m)) + (uint64_t)neg - own;
    const uint32_t w = (0x80000000u + (1u << SH) - 1u - (uint32_t)(d >> 32)) >> (SH + 1);
    return (4503599627370496.0 + (double)(1u << (BASE_LOG - 1))) - __hiloint2double(0x43300000, (int)w);
}
#endif


#if defined(__CUDA_ARCH__)
DEV uint32_t smem_addr_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
DEV void mbarrier_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
DEV void mbarrier_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
DEV void mbarrier_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity)
        : "memory");
}
// 8 complex doubles <-> 32 TMEM columns
DEV void tmem_store_cplx8(uint32_t taddr, cplx c0, cplx c1, cplx c2, cplx c3, cplx c4, cplx c5, cplx c6, cplx c7) {
    const cplx c[8] = {c0, c1, c2, c3, c4, c5, c6, c7};
    uint32_t v[32];
#pragma unroll
    for (int e = 0; e < 8; e++) {
        v[4 * e + 0] = (uint32_t)__double2loint(c[e].re);
        v[4 * e + 1] = (uint32_t)__double2hiint(c[e].re);
        v[4 * e + 2] = (uint32_t)__double2loint(c[e].im);
        v[4 * e + 3] = (uint32_t)__double2hiint(c[e].im);
    }
    tmem_st32(taddr, v);
}
DEV cplx cplx_from_words(const uint32_t (&v)[32], int e) {
    return cplx{__hiloint2double((int)v[4 * e + 1], (int)v[4 * e + 0]), __hiloint2double((int)v[4 * e + 3], (int)v[4 * e + 2])};
}
#endif

}  // namespace b200tfhe
