// fft_dev.cuh -- warp-level complex FFT of length N = 32*R, all butterflies in registers and
// warp shuffles (no shared-memory passes inside the transform).
//
// Replaces kissfft's mixed-radix kf_work (reference lib/kissfft/kiss_fft.c:236-300) for the two
// transform sizes on the hot path: real 320 -> complex 160 (R=5) and real 1024 -> complex 512
// (R=16).  Different algorithm, same mathematics: a Cooley-Tukey split N = R x 32 where
//   * lane l holds v[a] = z[l + 32a], a in [0,R)            (coalesced float2 loads)
//   * a length-R DFT runs over a inside each lane            (registers)
//   * twiddle w_N^(l*k1)
//   * a length-32 DFT runs across the lanes for every k1     (5 shuffle-xor DIF stages)
// Result: lane l, register k1 holds Z[k1 + R*bitrev5(l)].
//
// Twiddles come from a shared-memory table T[j] = exp(-2*pi*i*j/(2N)), j in [0,2N), computed on
// the host in double and rounded to float (same recipe as kiss_fft.c:361-367).
#pragma once

#include <cuda_runtime.h>

namespace fvad {

// Packed FP32 pairs (sm_100: FADD2 / FFMA2, one issue slot for both components of a complex value).  The spectral
// kernels are issue-bound, so complex add / subtract / scale-and-add go through these; each component is the same
// IEEE operation as the scalar form (a - b as fma(b, -1, a) is exact in the product), so results do not change.
// The complex multiply stays scalar: its packed form needs a swapped copy of an operand, which costs the slots back.
#ifndef FVAD_F32X2
#define FVAD_F32X2 1
#endif
#if FVAD_F32X2
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a); }
__device__ __forceinline__ float2 cfma(float2 a, float2 s, float2 c) { return __ffma2_rn(a, s, c); }
__device__ __forceinline__ float2 cscale(float2 a, float2 s) { return __fmul2_rn(a, s); }
#else
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cfma(float2 a, float2 s, float2 c) { return make_float2(fmaf(a.x, s.x, c.x), fmaf(a.y, s.y, c.y)); }
__device__ __forceinline__ float2 cscale(float2 a, float2 s) { return make_float2(a.x * s.x, a.y * s.y); }
#endif
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <bool INV>
__device__ __forceinline__ float2 tw(float2 w) { return INV ? make_float2(w.x, -w.y) : w; }
__device__ __forceinline__ int bitrev5(int l) { return (int)(__brev((unsigned)l) >> 27); }

// ---- in-register DFT of length R (forward: e^{-i...}; INV: e^{+i...}), unnormalised ----
template <int R, bool INV>
struct RegDft;

template <bool INV>
struct RegDft<1, INV> {
    static __device__ __forceinline__ void run(float2 (&)[1]) {}
};

// power-of-two R by recursive decimation in time; twiddles are literal constants so the
// trivial ones fold away at compile time.
template <int R, int K>
__device__ __forceinline__ float2 w_const() {
    // exp(-2*pi*i*K/R) for R in {2,4,8,16}
    constexpr float c16[16] = {1.0f, 0.92387953251128674f, 0.70710678118654752f, 0.38268343236508977f,
                               0.0f, -0.38268343236508977f, -0.70710678118654752f, -0.92387953251128674f,
                               -1.0f, -0.92387953251128674f, -0.70710678118654752f, -0.38268343236508977f,
                               0.0f, 0.38268343236508977f, 0.70710678118654752f, 0.92387953251128674f};
    constexpr int j = (K * (16 / R)) & 15;
    // sin(2*pi*j/16) = cos(2*pi*(j-4)/16)
    return make_float2(c16[j], -c16[(j + 12) & 15]);
}

template <int R, bool INV>
struct RegDft {
    static_assert(R == 2 || R == 4 || R == 8 || R == 16, "power-of-two register DFT");
    template <int K>
    static __device__ __forceinline__ void bfly(float2 (&v)[R], const float2 (&e)[R / 2], const float2 (&o)[R / 2]) {
        if constexpr (K < R / 2) {
            const float2 t = cmul(o[K], tw<INV>(w_const<R, K>()));
            v[K] = cadd(e[K], t);
            v[K + R / 2] = csub(e[K], t);
            bfly<K + 1>(v, e, o);
        }
    }
    static __device__ __forceinline__ void run(float2 (&v)[R]) {
        float2 e[R / 2], o[R / 2];
#pragma unroll
        for (int i = 0; i < R / 2; ++i) {
            e[i] = v[2 * i];
            o[i] = v[2 * i + 1];
        }
        RegDft<R / 2, INV>::run(e);
        RegDft<R / 2, INV>::run(o);
        bfly<0>(v, e, o);
    }
};

template <bool INV>
struct RegDft<5, INV> {
    static __device__ __forceinline__ void run(float2 (&v)[5]) {
        constexpr float c1 = 0.30901699437494742f, c2 = -0.80901699437494742f;  // cos(2pi/5), cos(4pi/5)
        constexpr float s1 = 0.95105651629515357f, s2 = 0.58778525229247313f;   // sin(2pi/5), sin(4pi/5)
        const float2 a1 = cadd(v[1], v[4]), d1 = csub(v[1], v[4]);
        const float2 a2 = cadd(v[2], v[3]), d2 = csub(v[2], v[3]);
        const float2 C1 = make_float2(c1, c1), C2 = make_float2(c2, c2), S1 = make_float2(s1, s1), S2 = make_float2(s2, s2);
        const float2 A = cfma(a2, C2, cfma(a1, C1, v[0]));
        const float2 B = cfma(a2, C1, cfma(a1, C2, v[0]));
        // P = s1*d1 + s2*d2 ; Q = s2*d1 - s1*d2 ; forward: Y1 = A - iP, Y4 = A + iP, Y2 = B - iQ, Y3 = B + iQ
        const float2 P = cfma(d2, S2, cscale(d1, S1));
        const float2 Q = cfma(d2, make_float2(-s1, -s1), cscale(d1, S2));
        // -i*(x+iy) = (y, -x)
        const float2 mP = INV ? make_float2(-P.y, P.x) : make_float2(P.y, -P.x);
        const float2 mQ = INV ? make_float2(-Q.y, Q.x) : make_float2(Q.y, -Q.x);
        v[0] = cadd(cadd(v[0], a1), a2);
        v[1] = cadd(A, mP);
        v[4] = csub(A, mP);
        v[2] = cadd(B, mQ);
        v[3] = csub(B, mQ);
    }
};

// Complex FFT over one warp.  T: shared-memory table of 2N entries (see header comment).
template <int R, bool INV>
__device__ __forceinline__ void warp_cfft(float2 (&v)[R], const float2 *__restrict__ T, int lane) {
    constexpr int N = 32 * R;
    RegDft<R, INV>::run(v);
#pragma unroll
    for (int k1 = 1; k1 < R; ++k1) v[k1] = cmul(v[k1], tw<INV>(T[2 * lane * k1]));
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
        const bool upper = (lane & h) != 0;
        const float2 w = tw<INV>(T[(lane & (h - 1)) * (N / h)]);
#pragma unroll
        for (int k1 = 0; k1 < R; ++k1) {
            float2 b;
            b.x = __shfl_xor_sync(0xffffffffu, v[k1].x, h);
            b.y = __shfl_xor_sync(0xffffffffu, v[k1].y, h);
            v[k1] = upper ? cmul(csub(b, v[k1]), w) : cadd(v[k1], b);
        }
    }
}

// Same transform with every lane-constant twiddle hoisted into registers (a kernel that runs many
// frames per warp initialises this once).  The half-select of each butterfly is folded into the
// operands: a = sg*v + b (sg = -1 in the upper half: b - v, +1 in the lower: v + b), then a * ws with
// ws = (1, 0) in the lower half -- bit-identical to warp_cfft, ~40 % fewer instructions.
template <int R, bool INV>
struct WarpFftTw {
    float2 t1[R > 1 ? R - 1 : 1];  // w_N^(lane*k1), k1 = 1..R-1
    float2 ws[4];                  // stages h = 16, 8, 4, 2
    float sg[5];                   // stages h = 16, 8, 4, 2, 1
    __device__ __forceinline__ void init(const float2 *__restrict__ T, int lane) {
        constexpr int N = 32 * R;
#pragma unroll
        for (int k1 = 1; k1 < R; ++k1) t1[k1 - 1] = tw<INV>(T[2 * lane * k1]);
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const int h = 16 >> i;
            const bool upper = (lane & h) != 0;
            sg[i] = upper ? -1.0f : 1.0f;
            if (i < 4) ws[i] = upper ? tw<INV>(T[(lane & (h - 1)) * (N / h)]) : make_float2(1.0f, 0.0f);
        }
    }
};

template <int R, bool INV>
__device__ __forceinline__ void warp_cfft_tw(float2 (&v)[R], const WarpFftTw<R, INV> &w) {
    RegDft<R, INV>::run(v);
#pragma unroll
    for (int k1 = 1; k1 < R; ++k1) v[k1] = cmul(v[k1], w.t1[k1 - 1]);
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const int h = 16 >> i;
#pragma unroll
        for (int k1 = 0; k1 < R; ++k1) {
            float2 b;
            b.x = __shfl_xor_sync(0xffffffffu, v[k1].x, h);
            b.y = __shfl_xor_sync(0xffffffffu, v[k1].y, h);
            const float2 a = cfma(v[k1], make_float2(w.sg[i], w.sg[i]), b);
            v[k1] = i < 4 ? cmul(a, w.ws[i]) : a;  // the last stage's twiddle is 1 for every lane
        }
    }
}

// index of the value held by (lane, register k1) after warp_cfft
template <int R>
__device__ __forceinline__ int cfft_out_index(int lane, int k1) { return k1 + R * bitrev5(lane); }

// Real-FFT unpack (what kiss_fftr.c:93-116 does): bin k of the length-2N real transform from the
// complex transform Z of the packed signal.  k in [0,N]; Z indices are taken mod N.
__device__ __forceinline__ float2 rfft_bin(const float2 *__restrict__ Z, const float2 *__restrict__ T, int N, int k) {
    const float2 a = Z[k == N ? 0 : k];
    const float2 bb = Z[k == 0 ? 0 : N - k];
    const float2 b = make_float2(bb.x, -bb.y);
    const float2 e = cadd(a, b), d = csub(a, b);
    // o = d * (-i * T[k])   (kissfft's super twiddle exp(-i*pi*(k/N + 1/2)))
    const float2 t = T[k];
    const float2 mi_t = make_float2(t.y, -t.x);  // -i*(x+iy) = (y,-x)
    const float2 o = cmul(d, mi_t);
    // DC and Nyquist are purely real (kiss_fftr.c:82-89 writes their imaginary part as 0)
    return make_float2((e.x + o.x) * 0.5f, (k == 0 || k == N) ? 0.0f : (e.y + o.y) * 0.5f);
}

// The same unpack in kissfft's own pair form (kiss_fftr.c:93-116): bins k and N-k from Z[k], Z[N-k] and ONE
// super twiddle st = exp(-i*pi*(k/N + 1/2)) = -i*T[k]; k in [1, N/2].
__device__ __forceinline__ void rfft_pair(float2 zk, float2 znk, float2 st, float2 &Xk, float2 &Xnk) {
    const float2 f1 = make_float2(zk.x + znk.x, zk.y - znk.y);
    const float2 f2 = make_float2(zk.x - znk.x, zk.y + znk.y);
    const float2 t = cmul(f2, st);
    Xk = make_float2((f1.x + t.x) * 0.5f, (f1.y + t.y) * 0.5f);
    Xnk = make_float2((f1.x - t.x) * 0.5f, (t.y - f1.y) * 0.5f);
}

// rfft_pair for a transform whose input was pre-scaled by 1/2 (a kernel folds the 0.5 into its window: scaling
// by a power of two commutes with every rounding, so the bins are bit-identical and four multiplies are saved).
__device__ __forceinline__ void rfft_pair_halved(float2 zk, float2 znk, float2 st, float2 &Xk, float2 &Xnk) {
    const float2 f1 = cfma(znk, make_float2(1.0f, -1.0f), zk);  // (zk.x + znk.x, zk.y - znk.y), exact products
    const float2 f2 = cfma(znk, make_float2(-1.0f, 1.0f), zk);  // (zk.x - znk.x, zk.y + znk.y)
    const float2 t = cmul(f2, st);
    Xk = cadd(f1, t);
    Xnk = make_float2(f1.x - t.x, t.y - f1.y);
}

// Inverse pre-pack (kiss_fftri, kiss_fftr.c:131-153): Zt[k] = (X[k]+conj(X[N-k])) + i*conj(T[k])*(X[k]-conj(X[N-k]))
__device__ __forceinline__ float2 irfft_pack(const float2 *__restrict__ X, const float2 *__restrict__ T, int N, int k) {
    const float2 a = X[k];
    const float2 bb = X[N - k];
    const float2 b = make_float2(bb.x, -bb.y);
    const float2 e = cadd(a, b), d = csub(a, b);
    const float2 t = T[k];                       // (c, s) = exp(-i*pi*k/N)  -> conj = (c, -s)
    const float2 it = make_float2(t.y, t.x);     // i*(c - i s) = (s, c)
    return cadd(e, cmul(d, it));
}

}  // namespace fvad
