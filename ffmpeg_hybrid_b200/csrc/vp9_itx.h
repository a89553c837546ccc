/*
 * vp9_itx.h -- VP9 inverse transforms (DCT / ADST 4..32, WHT 4) in exact integer arithmetic.
 *
 * Semantics follow the reference's itxfm_add slots, libavcodec/vp9dsp_template.c:1155-1778:
 *   - every rotation is (a*c0 +- b*c1 + 8192) >> 14 on 14-bit cosine constants;
 *   - 8-bit streams: all arithmetic is 32-bit two's-complement WRAPPING (the reference's
 *     `dctint` is int and it multiplies by unsigned constants, :1325-1340), values are
 *     truncated to int16_t when stored between the passes (`dctcoef tmp[]`, :1162,1181);
 *   - 10/12-bit streams: 64-bit arithmetic, int32_t storage (vp9dsp_10bpp.c:24,
 *     bit_depth_template.c:82);
 *   - the first pass runs along block[i + x*N] (x = 0..N-1) for each i, the second pass along
 *     tmp[x*N + i] and lands in pixel column i (:1180-1190);
 *   - output: dst = clip(dst + ((int)(out + (1u << (bits-1))) >> bits)), bits = 4/5/6/6/0;
 *   - DCT_DCT blocks with eob == 1 take the DC-only shortcut (:1165-1178).
 *
 * The formulation here is recursive (an N-point DCT is an N/2-point DCT of the even inputs
 * plus an odd-input butterfly network), shared by the CUDA kernels and by the host model in
 * oracle/; it is written for registers: in[] and out[] are small local arrays.
 */
#ifndef VP9_ITX_H
#define VP9_ITX_H

#include <stdint.h>

#ifdef __CUDACC__
#define VP9_HD __host__ __device__ __forceinline__
#define VP9_HDM __host__ __device__ __forceinline__
#else
#define VP9_HD static inline __attribute__((always_inline))
#define VP9_HDM inline __attribute__((always_inline))
#endif

namespace vp9 {

/* 32-bit wrapping integer: what `int` is in the reference's 8-bit instantiation under -fwrapv */
struct W32 {
    int32_t v;
    VP9_HDM W32() : v(0) {}
    VP9_HDM explicit W32(int64_t x) : v((int32_t)(uint32_t)(uint64_t)x) {}
    VP9_HDM W32 operator+(W32 o) const { return W32((int64_t)(int32_t)((uint32_t)v + (uint32_t)o.v)); }
    VP9_HDM W32 operator-(W32 o) const { return W32((int64_t)(int32_t)((uint32_t)v - (uint32_t)o.v)); }
    VP9_HDM W32 operator-() const { return W32((int64_t)(int32_t)(0u - (uint32_t)v)); }
    VP9_HDM W32 operator*(int c) const { return W32((int64_t)(int32_t)((uint32_t)v * (uint32_t)c)); }
    VP9_HDM W32 rs14() const { return W32((int64_t)((int32_t)((uint32_t)v + 8192u) >> 14)); }
    VP9_HDM W32 sra(int n) const { return W32((int64_t)(v >> n)); }
};
/* 64-bit integer: `dctint` of the 10/12-bit instantiation */
struct W64 {
    int64_t v;
    VP9_HDM W64() : v(0) {}
    VP9_HDM explicit W64(int64_t x) : v(x) {}
    VP9_HDM W64 operator+(W64 o) const { return W64((int64_t)((uint64_t)v + (uint64_t)o.v)); }
    VP9_HDM W64 operator-(W64 o) const { return W64((int64_t)((uint64_t)v - (uint64_t)o.v)); }
    VP9_HDM W64 operator-() const { return W64((int64_t)(0ull - (uint64_t)v)); }
    VP9_HDM W64 operator*(int c) const { return W64((int64_t)((uint64_t)v * (uint64_t)(int64_t)c)); }
    VP9_HDM W64 rs14() const { return W64((int64_t)((uint64_t)v + 8192ull) >> 14); }
    VP9_HDM W64 sra(int n) const { return W64(v >> n); }
};

/* (a*c0 - b*c1 + 8192) >> 14 and (a*c1 + b*c0 + 8192) >> 14 */
template <class T> VP9_HD void rot(T a, T b, int c0, int c1, T &lo, T &hi)
{
    lo = (a * c0 - b * c1).rs14();
    hi = (a * c1 + b * c0).rs14();
}
/* ((a - b) * 11585 + 8192) >> 14 and ((a + b) * 11585 + 8192) >> 14 */
template <class T> VP9_HD void half(T a, T b, T &diff, T &sum)
{
    diff = ((a - b) * 11585).rs14();
    sum  = ((a + b) * 11585).rs14();
}

/* ---- DCT ---------------------------------------------------------------------------------
 * in[] is read with stride `s` (so the recursion can pick every other input);
 * out[0..N-1] natural order. */
template <class T> VP9_HD void idct4(const T *in, int s, T *out)
{
    T e0, e1, o0, o1;
    half(in[0], in[2 * s], e1, e0);               /* e0 = (in0+in2)c, e1 = (in0-in2)c */
    rot(in[s], in[3 * s], 6270, 15137, o0, o1);   /* o0 = in1*6270 - in3*15137 */
    out[0] = e0 + o1;
    out[1] = e1 + o0;
    out[2] = e1 - o0;
    out[3] = e0 - o1;
}

template <class T> VP9_HD void idct8(const T *in, int s, T *out)
{
    T e[4], a, b, c, d;
    idct4(in, 2 * s, e);
    rot(in[s], in[7 * s], 3196, 16069, a, d);      /* a=t4a d=t7a */
    rot(in[5 * s], in[3 * s], 13623, 9102, b, c);  /* b=t5a c=t6a */
    T t4 = a + b, t5 = a - b, t7 = d + c, t6 = d - c;
    half(t6, t5, b, c);                            /* b = (t6-t5)c, c = (t6+t5)c */
    T o[4] = { t7, c, b, t4 };
#pragma unroll
    for (int k = 0; k < 4; k++) {
        out[k]     = e[k] + o[k];
        out[7 - k] = e[k] - o[k];
    }
}

template <class T> VP9_HD void idct16(const T *in, int s, T *out)
{
    T e[8], u[8];   /* u[k] stands for t(8+k) */
    idct8(in, 2 * s, e);
    rot(in[1 * s],  in[15 * s],  1606, 16305, u[0], u[7]);
    rot(in[9 * s],  in[7 * s],  12665, 10394, u[1], u[6]);
    rot(in[5 * s],  in[11 * s],  7723, 14449, u[2], u[5]);
    rot(in[13 * s], in[3 * s],  15679,  4756, u[3], u[4]);
    T t8 = u[0] + u[1], t9 = u[0] - u[1], t10 = u[3] - u[2], t11 = u[3] + u[2];
    T t12 = u[4] + u[5], t13 = u[4] - u[5], t14 = u[7] - u[6], t15 = u[7] + u[6];
    T t9a, t14a, t10a, t13a;
    rot(t14, t9, 6270, 15137, t9a, t14a);
    t10a = (-(t13 * 15137 + t10 * 6270)).rs14();
    t13a = (t13 * 6270 - t10 * 15137).rs14();
    T t8a = t8 + t11, t11a = t8 - t11, t12a = t15 - t12, t15a = t15 + t12;
    t9 = t9a + t10a;  t10 = t9a - t10a;
    t13 = t14a - t13a; t14 = t14a + t13a;
    half(t13, t10, t10a, t13a);
    half(t12a, t11a, t11, t12);
    T o[8] = { t15a, t14, t13a, t12, t11, t10a, t9, t8a };
#pragma unroll
    for (int k = 0; k < 8; k++) {
        out[k]      = e[k] + o[k];
        out[15 - k] = e[k] - o[k];
    }
}

template <class T> VP9_HD void idct32(const T *in, int s, T *out)
{
    T e[16];
    idct16(in, 2 * s, e);
    T a16, a17, a18, a19, a20, a21, a22, a23, a24, a25, a26, a27, a28, a29, a30, a31;
    rot(in[1 * s],  in[31 * s],   804, 16364, a16, a31);
    rot(in[17 * s], in[15 * s], 12140, 11003, a17, a30);
    rot(in[9 * s],  in[23 * s],  7005, 14811, a18, a29);
    rot(in[25 * s], in[7 * s],  15426,  5520, a19, a28);
    rot(in[5 * s],  in[27 * s],  3981, 15893, a20, a27);
    rot(in[21 * s], in[11 * s], 14053,  8423, a21, a26);
    rot(in[13 * s], in[19 * s],  9760, 13160, a22, a25);
    rot(in[29 * s], in[3 * s],  16207,  2404, a23, a24);

    T t16 = a16 + a17, t17 = a16 - a17, t18 = a19 - a18, t19 = a19 + a18;
    T t20 = a20 + a21, t21 = a20 - a21, t22 = a23 - a22, t23 = a23 + a22;
    T t24 = a24 + a25, t25 = a24 - a25, t26 = a27 - a26, t27 = a27 + a26;
    T t28 = a28 + a29, t29 = a28 - a29, t30 = a31 - a30, t31 = a31 + a30;

    rot(t30, t17, 3196, 16069, a17, a30);
    a18 = (-(t29 * 16069 + t18 * 3196)).rs14();
    a29 = (t29 * 3196 - t18 * 16069).rs14();
    rot(t26, t21, 13623, 9102, a21, a26);
    a22 = (-(t25 * 9102 + t22 * 13623)).rs14();
    a25 = (t25 * 13623 - t22 * 9102).rs14();

    a16 = t16 + t19;  t17 = a17 + a18;  t18 = a17 - a18;  a19 = t16 - t19;
    a20 = t23 - t20;  t21 = a22 - a21;  t22 = a22 + a21;  a23 = t23 + t20;
    a24 = t24 + t27;  t25 = a25 + a26;  t26 = a25 - a26;  a27 = t24 - t27;
    a28 = t31 - t28;  t29 = a30 - a29;  t30 = a30 + a29;  a31 = t31 + t28;

    rot(t29, t18, 6270, 15137, a18, a29);
    rot(a28, a19, 6270, 15137, t19, t28);
    t20 = (-(a27 * 15137 + a20 * 6270)).rs14();
    t27 = (a27 * 6270 - a20 * 15137).rs14();
    a21 = (-(t26 * 15137 + t21 * 6270)).rs14();
    a26 = (t26 * 6270 - t21 * 15137).rs14();

    t16 = a16 + a23;  a17 = t17 + t22;  t18 = a18 + a21;  a19 = t19 + t20;
    a20 = t19 - t20;  t21 = a18 - a21;  a22 = t17 - t22;  t23 = a16 - a23;
    t24 = a31 - a24;  a25 = t30 - t25;  t26 = a29 - a26;  a27 = t28 - t27;
    a28 = t28 + t27;  t29 = a29 + a26;  a30 = t30 + t25;  t31 = a31 + a24;

    half(a27, a20, t20, t27);
    half(t26, t21, a21, a26);
    half(a25, a22, t22, t25);
    half(t24, t23, a23, a24);

    T o[16] = { t31, a30, t29, a28, t27, a26, t25, a24, a23, t22, a21, t20, a19, t18, a17, t16 };
#pragma unroll
    for (int k = 0; k < 16; k++) {
        out[k]      = e[k] + o[k];
        out[31 - k] = e[k] - o[k];
    }
}

/* ---- ADST (vp9dsp_template.c:1217-1232, 1273-1314, 1408-1507) -------------------------------- */
template <class T> VP9_HD void iadst4(const T *in, int s, T *out)
{
    T x0 = in[0], x1 = in[s], x2 = in[2 * s], x3 = in[3 * s];
    T t0 = x0 * 5283 + x2 * 15212 + x3 * 9929;
    T t1 = x0 * 9929 - x2 * 5283 - x3 * 15212;
    T t2 = (x0 - x2 + x3) * 13377;
    T t3 = x1 * 13377;
    out[0] = (t0 + t3).rs14();
    out[1] = (t1 + t3).rs14();
    out[2] = t2.rs14();
    out[3] = (t0 + t1 - t3).rs14();
}

template <class T> VP9_HD void iadst8(const T *in, int s, T *out)
{
    T p0 = in[7 * s] * 16305 + in[0] * 1606,      p1 = in[7 * s] * 1606 - in[0] * 16305;
    T p2 = in[5 * s] * 14449 + in[2 * s] * 7723,  p3 = in[5 * s] * 7723 - in[2 * s] * 14449;
    T p4 = in[3 * s] * 10394 + in[4 * s] * 12665, p5 = in[3 * s] * 12665 - in[4 * s] * 10394;
    T p6 = in[1 * s] * 4756 + in[6 * s] * 15679,  p7 = in[1 * s] * 15679 - in[6 * s] * 4756;
    T t0 = (p0 + p4).rs14(), t1 = (p1 + p5).rs14(), t2 = (p2 + p6).rs14(), t3 = (p3 + p7).rs14();
    T t4 = (p0 - p4).rs14(), t5 = (p1 - p5).rs14(), t6 = (p2 - p6).rs14(), t7 = (p3 - p7).rs14();
    p4 = t4 * 15137 + t5 * 6270;  p5 = t4 * 6270 - t5 * 15137;
    p6 = t7 * 15137 - t6 * 6270;  p7 = t7 * 6270 + t6 * 15137;
    out[0] = t0 + t2;
    out[7] = -(t1 + t3);
    t2 = t0 - t2;
    t3 = t1 - t3;
    out[1] = -((p4 + p6).rs14());
    out[6] = (p5 + p7).rs14();
    t6 = (p4 - p6).rs14();
    t7 = (p5 - p7).rs14();
    out[3] = -(((t2 + t3) * 11585).rs14());
    out[4] = ((t2 - t3) * 11585).rs14();
    out[2] = ((t6 + t7) * 11585).rs14();
    out[5] = -(((t6 - t7) * 11585).rs14());
}

template <class T> VP9_HD void iadst16(const T *in, int s, T *out)
{
    const int c0[8] = { 16364, 15893, 14811, 13160, 11003, 8423, 5520, 2404 };
    const int c1[8] = { 804, 3981, 7005, 9760, 12140, 14053, 15426, 16207 };
    T p[16], q[16];
    /* pair k: hi input 15-2k, lo input 2k */
#pragma unroll
    for (int k = 0; k < 8; k++) {
        T hi = in[(15 - 2 * k) * s], lo = in[(2 * k) * s];
        p[2 * k]     = hi * c0[k] + lo * c1[k];
        p[2 * k + 1] = hi * c1[k] - lo * c0[k];
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        q[k]     = (p[k] + p[k + 8]).rs14();
        q[k + 8] = (p[k] - p[k + 8]).rs14();
    }
    T t8  = q[8] * 16069 + q[9] * 3196,    t9  = q[8] * 3196 - q[9] * 16069;
    T t10 = q[10] * 9102 + q[11] * 13623,  t11 = q[10] * 13623 - q[11] * 9102;
    T t12 = q[13] * 16069 - q[12] * 3196,  t13 = q[13] * 3196 + q[12] * 16069;
    T t14 = q[15] * 9102 - q[14] * 13623,  t15 = q[15] * 13623 + q[14] * 9102;
    T t0 = q[0] + q[4], t1 = q[1] + q[5], t2 = q[2] + q[6], t3 = q[3] + q[7];
    T t4 = q[0] - q[4], t5 = q[1] - q[5], t6 = q[2] - q[6], t7 = q[3] - q[7];
    T t8a  = (t8 + t12).rs14(),  t9a  = (t9 + t13).rs14();
    T t10a = (t10 + t14).rs14(), t11a = (t11 + t15).rs14();
    T t12a = (t8 - t12).rs14(),  t13a = (t9 - t13).rs14();
    T t14a = (t10 - t14).rs14(), t15a = (t11 - t15).rs14();
    T t4a = t4 * 15137 + t5 * 6270,  t5a = t4 * 6270 - t5 * 15137;
    T t6a = t7 * 15137 - t6 * 6270,  t7a = t7 * 6270 + t6 * 15137;
    t12 = t12a * 15137 + t13a * 6270;  t13 = t12a * 6270 - t13a * 15137;
    t14 = t15a * 15137 - t14a * 6270;  t15 = t15a * 6270 + t14a * 15137;

    out[0]  = t0 + t2;
    out[15] = -(t1 + t3);
    T t2a = t0 - t2, t3a = t1 - t3;
    out[3]  = -((t4a + t6a).rs14());
    out[12] = (t5a + t7a).rs14();
    t6 = (t4a - t6a).rs14();
    t7 = (t5a - t7a).rs14();
    out[1]  = -(t8a + t10a);
    out[14] = t9a + t11a;
    t10 = t8a - t10a;
    t11 = t9a - t11a;
    out[2]  = (t12 + t14).rs14();
    out[13] = -((t13 + t15).rs14());
    t14a = (t12 - t14).rs14();
    t15a = (t13 - t15).rs14();
    out[7]  = ((-(t2a + t3a)) * 11585).rs14();
    out[8]  = ((t2a - t3a) * 11585).rs14();
    out[4]  = ((t7 + t6) * 11585).rs14();
    out[11] = ((t7 - t6) * 11585).rs14();
    out[6]  = ((t11 + t10) * 11585).rs14();
    out[9]  = ((t11 - t10) * 11585).rs14();
    out[5]  = ((-(t14a + t15a)) * 11585).rs14();
    out[10] = ((t14a - t15a) * 11585).rs14();
}

/* ---- WHT (lossless), vp9dsp_template.c:1719-1748: always 32-bit int temporaries ------------- */
VP9_HD void iwht4(const W32 *in, int s, W32 *out, int pass)
{
    W32 t0 = in[0], t1 = in[3 * s], t2 = in[s], t3 = in[2 * s], t4;
    if (pass == 0) {
        t0 = t0.sra(2); t1 = t1.sra(2); t2 = t2.sra(2); t3 = t3.sra(2);
    }
    t0 = t0 + t2;
    t3 = t3 - t1;
    t4 = (t0 - t3).sra(1);
    t1 = t4 - t1;
    t2 = t4 - t2;
    t0 = t0 - t1;
    t3 = t3 + t2;
    out[0] = t0; out[1] = t1; out[2] = t2; out[3] = t3;
}

/* One 1-D transform of length N (N = 4 << tx), adst selects ADST (never for N == 32). */
template <class T, int N> VP9_HD void itx_1d(const T *in, T *out, bool adst)
{
    if constexpr (N == 4)       { if (adst) iadst4(in, 1, out);  else idct4(in, 1, out); }
    else if constexpr (N == 8)  { if (adst) iadst8(in, 1, out);  else idct8(in, 1, out); }
    else if constexpr (N == 16) { if (adst) iadst16(in, 1, out); else idct16(in, 1, out); }
    else                        { idct32(in, 1, out); }
}

VP9_HD int clip_pixel(int v, int bpp)
{
    int mx = (1 << bpp) - 1;
    return v < 0 ? 0 : (v > mx ? mx : v);
}

/* residual sample from a second-pass output as stored in dctcoef (int16_t or int32_t):
 * (int)(out + (1u << (bits-1))) >> bits, bits = 0 means identity (vp9dsp_template.c:1186-1189) */
VP9_HD int itx_round(int32_t stored, int bits)
{
    if (!bits)
        return stored;
    return (int32_t)((uint32_t)stored + (1u << (bits - 1))) >> bits;
}

/* rounding shift of the 2-D transform of size tx (0..3), lossless WHT = 0 */
VP9_HD int itx_bits(int tx, bool lossless)
{
    return lossless ? 0 : (tx == 0 ? 4 : tx == 1 ? 5 : 6);
}

/* DC-only shortcut value `t` (before the final rounding shift), vp9dsp_template.c:1166-1167 */
template <bool HBD> VP9_HD int32_t itx_dc_only(int32_t dc)
{
    if (HBD) {
        int64_t t = (((int64_t)dc * 11585 + (1 << 13)) >> 14);
        t = (t * 11585 + (1 << 13)) >> 14;
        return (int32_t)(uint32_t)(uint64_t)t;
    } else {
        W32 t = (W32(dc) * 11585).rs14();
        t = (t * 11585).rs14();
        return t.v;
    }
}

} /* namespace vp9 */
#endif /* VP9_ITX_H */
