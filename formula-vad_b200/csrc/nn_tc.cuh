// nn_tc.cuh -- kernel argument blocks of the tcgen05 NSNet2 backend (nn_tc.cu, nn_tc_gemm.cu, nn_tc_gru.cu).
#pragma once
#include "common.cuh"

#include <cuda.h>
#include <cuda_fp16.h>

namespace fvad {
namespace tc {

constexpr int kTileM = 128;            // rows (sequences) per M-tile = TMEM lanes
constexpr int kMaxNTiles = 13;

// bytes of one activation tile of `cols` columns (cols % 16 == 0) in the core-matrix tiled layout
__host__ __device__ constexpr size_t tile_bytes(int cols) { return (size_t)kTileM * cols * 2; }
// byte offset of element (row, col) inside a tile of R rows
__host__ __device__ constexpr size_t tiled_off(int row, int col, int R) {
    return (size_t)(col >> 3) * (R * 16) + (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16 + (size_t)(col & 7) * 2;
}

// C = act(A * B^T + bias) over M-tiles of 128 rows.
//   A: activation tiles, K columns.  M-tile mt -> (q = mt / T, t = mt % T); A tile index = q*A_q + t + A_t0.
//   B: weight image: for N-tile j (nt_size[j] rows) a contiguous block at B + nt_off[j] bytes, tiled with R = nt_size[j].
//   out_mode 0: C tiles of Nc columns (fp16), tile index q*C_q + t + C_t0, N-tile j at column nt_col[j].
//   out_mode 1: gains scatter (f32 row-major [sc][50K][161]), frame-in-chunk = t + C_t0, columns < n_valid.
struct GemmArgs {
    const __half *A;
    const __half *B;
    const __half *B2;   // CTA-pair image: per N-tile [rows 0..NT/2 | rows NT/2..NT], each half tiled with R = NT/2
    const float *bias;  // [Nc]
    __half *C;
    float *gains;
    int K, Nc, n_valid;
    int n_tiles;
    int nt_size[kMaxNTiles], nt_col[kMaxNTiles];
    unsigned nt_off[kMaxNTiles];
    int mt_count, T, A_q, A_t0, C_q, C_t0;
    int act;       // 0 none, 1 relu, 2 sigmoid
    int out_mode;
    unsigned long long *trace;  // debug (FVAD_GEMM_TRACE): CTA 0 logs clock64 for its first 4 M-tiles, nullptr otherwise
    int a_tmem;    // 1: the A tile lives in TMEM (tcgen05.cp + A-from-TMEM MMAs), all shared memory is weight ring; needs
                   //    K/2 + 2*max(nt_size) <= 512 TMEM columns
    int dbg;       // experiments only (FVAD_GEMM_DBG; results are then wrong): bit 0 epilogue does no work, bit 1 producer copies nothing,
                   // bit 2 (pair kernel) no C stores
    int seq0, n_seq, Kchunks;  // scatter / gather: first global sequence of tile q = 0, sequences in total, chunks per stream-channel
    const __half *gather_rows; // pair kernel, K = 176 only: A is not tiles but the fp16 feature rows [SC][4+50*Kchunks][176]; row r of
                               // M-tile (q, t) is row 50*j + t of stream-channel sc, (sc, j) = divmod(seq0 + 128 q + r, Kchunks) --
                               // the 54-row windows are gathered by the A stagers themselves (no pack pass); nullptr otherwise
};
void launch_tc_gemm(const GemmArgs &a, int n_sm, cudaStream_t s);

// CTA-pair version (nn_tc_gemm2.cu; cta_group::2, M = 256, A in TMEM).  The weights come from the pair image B2 through
// TMA: one tensor map per (N-tile, half), passed to the kernel as a __grid_constant__ parameter.
struct PairMaps {
    CUtensorMap m[2 * kMaxNTiles];
};
void encode_tiled_map(CUtensorMap *m, CUtensorMapDataType dt, unsigned rank, void *base, const cuuint64_t *gdim, const cuuint64_t *gstr,
                      const cuuint32_t *box);
int pair_slab_ksteps(int K);
void build_pair_maps(PairMaps &maps, const void *img2, int K, int n_tiles, const int *nt_size, const unsigned *nt_off);
void launch_tc_gemm2(const GemmArgs &a, const PairMaps &maps, int n_sm, cudaStream_t s);

// One GRU layer over 54 steps for Q tiles of 128 sequences.
//   xp:   tiles [q][54] of 1200 columns: W x_t + Wb (+Rb for z,r), columns in chunk-interleaved order
//         (chunk c = units 16c..16c+15: [z(16) | r(16) | n(16)])
//   Rimg: 25 chunk images (48 rows x 400 K, tiled with R = 48), same row order
//   rbn:  Rb of the n gate, natural unit order [400]
//   hout: tiles [q][54] of 400 columns (steps < t_store_min are not written)
struct GruArgs {
    const __half *xp;
    const __half *Rimg;
    const float *rbn;
    __half *hout;
    int Q, t_store_min;
    unsigned long long *trace;  // debug (FVAD_GRU_TRACE): CTA 0 logs clock64 at [step][chunk][event], nullptr otherwise
    int dbg = 0;                // experiments only (FVAD_GRU_DBG, pair kernel; results are then WRONG, only the time means something):
                                // 1 no accumulator/h loads from TMEM, 2 no weight copies, 4 no x-projection loads, 8 no h_t stores to HBM,
                                // 16 no gate arithmetic, 32 gates without MUFU.TANH, 64 MUFU.TANH.F16x2 (profiles/r02_gru_knockouts.txt)
};
void launch_tc_gru(const GruArgs &a, int n_sm, cudaStream_t s);
// CTA-pair version (nn_tc_gru2.cu): Rimg is the split image (per chunk two 24-row halves tiled with R = 24), fetched by TMA
// through `rmap` (build_gru_pair_map)
void build_gru_pair_map(CUtensorMap &rmap, const void *Rimg2);
void launch_tc_gru2(const GruArgs &a, const CUtensorMap &rmap, int n_sm, cudaStream_t s);

// First GRU layer with the (fc1-folded) input projection fused in (nn_tc_gru3.cu): reads the fp16 feature rows directly.
//   feat16: [SC][4+50*Kchunks][176]; row r of tile q is sequence seq0 + 128 q + r = (sc, j) by divmod Kchunks, its 54-row
//           window starts at row 50 j of stream-channel sc
//   Rimg3 / Wimg3 (through the tensor maps of build_gru3_maps): per chunk two 24-row halves; R rows in [n z r] order over
//           K = 400, folded W rows in [z r n] order over K = 176 with the bias as an fp16 hi + lo pair in columns 161, 162
struct Gru3Args {
    const __half *feat16;
    const float *rbn;
    __half *hout;
    int Q, t_store_min, seq0, n_seq, Kchunks;
};
void build_gru3_maps(CUtensorMap &rmap, CUtensorMap &wmap, const void *Rimg3, const void *Wimg3);
void launch_tc_gru3(const Gru3Args &a, const CUtensorMap &rmap, const CUtensorMap &wmap, int n_sm, cudaStream_t s);

// feature rows fp16 [SC][4+50K][176] -> fp16 tiles [q][54] of 176 columns for sequences seq0 .. seq0+128*Q-1
struct PackArgs {
    const __half *feat16;
    __half *A0;
    int Q, seq0, n_seq, Kchunks;
};
void launch_tc_pack(const PackArgs &a, cudaStream_t s);

}  // namespace tc
}  // namespace fvad
