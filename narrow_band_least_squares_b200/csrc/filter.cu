// filter.cu — K1: per-band IIR band-pass as cascaded biquads over every channel x band, with a
// time-chunked scan of the linear-recurrence state, and the Hann taper fused into the last store.
//
// Replaces helpers.py:108-141 filter_data of the reference (obspy Stream.filter('bandpass',
// zerophase=True) = scipy sosfilt forward, then sosfilt over the reversed output, zero state
// both times, no padding; then Trace.taper(0.01)).  The recurrence is scipy's direct-form-II
// transposed section loop with the SAME unfused multiply/add order, so a chunk
// that starts from the true state reproduces scipy bit for bit.
//
// Parallel structure (per direction):
//   phase 1  every chunk of L samples runs from zero state over its last K_b samples (all of
//            them in exact mode) and stores its end state                    -> reads only
//   phase 2  one thread per (band, channel) carries states across chunks:
//            s_in[k+1] = A_b^L s_in[k] + e[k]          (A_b^L from the host, FP64)
//   phase 3  every chunk reruns from its true state and writes its output  -> read + write
// A warp owns 32 consecutive chunks of one (band, channel); lanes walk their chunks in
// lockstep, so each step moves a [32 chunks x 16 samples] tile: cp.async (16 B per lane,
// eight lanes per 128-B line) into a padded shared-memory tile, lane r computes row r, and the
// results leave through the same tile as full 128-B lines.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <complex>
#include <type_traits>
#include <vector>

#include "common.cuh"

namespace nbls {

constexpr int FT_TILE = 32;    // samples per chunk per step (one 128-B line; 32 = 256 B was measured: no gain)
constexpr int FT_ROW = FT_TILE + 2;   // padded row stride in doubles: (8 TILE + 16) B keeps 16-B alignment and
                               // makes lane r's LDS.128 hit banks 4r..4r+3 (conflict-free)
constexpr int FT_STAGES = 3;   // cp.async ring depth
constexpr int FT_WARPS = 4;    // bands per block
constexpr int FT_TILE_DOUBLES = 32 * FT_ROW;
constexpr int FT_SEGS = FT_TILE / 2;          // 16-byte segments per tile row
constexpr int FT_ROWS_PER_REQ = 32 / FT_SEGS; // rows covered by one warp-wide request
constexpr int FT_REQS = 32 / FT_ROWS_PER_REQ; // requests per tile

struct FilterArgs {
    const double* in;
    int64_t in_row, in_band;   // in_band == 0: every band reads the same rows (the raw record)
    double* out;               // nullptr: phase 1 (states only)
    int64_t out_row, out_band;
    const double* coef;        // [B][S][5] b0 b1 b2 a1 a2
    const int* proc_tiles;     // [B] tiles per chunk to run in phase 1 (from the chunk's end)
    const double* state_in;    // [B][C][sin_n][2S] or nullptr (zero); chunk k starts from entry sin_mul * k + sin_off
    double* state_out;         // [B][C][nchunks][2S] or nullptr
    int sin_n, sin_mul, sin_off;   // the state pass may run on SUB-chunks (forward pass): sin_n = its chunk count
    int B, C;
    int64_t T;
    int L, nchunks;
    const double* taper_up;
    const double* taper_down;
    int64_t taper_len, g0, G;
    int nsplit;                // biquad_state_shared_kernel: blockIdx.z = g takes bands g, g + nsplit, g + 2 nsplit, ...
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// One warp, 32 consecutive chunks of one (band, channel).  MODE 0 is the interior fast path: every chunk of the warp is
// whole, so loads and stores carry no bounds or taper logic (the Hann ramps of whole chunks are multiplied in afterwards by
// taper_rmw_kernel).  MODE 2: the ragged record end (or a single-chunk run): bounds checks, taper fused into the store.
template <int S, bool REV, int MODE, bool STATE_ONLY, bool FUSED>
__device__ __forceinline__ void biquad_warp(const FilterArgs& a, double* ring, int band, int chan, int64_t chunk0,
                                            int lane) {
    constexpr bool EDGE = MODE == 2;
    const int ntiles = a.L / FT_TILE;
    int first = 0;
    if (a.proc_tiles) first = ntiles - min(ntiles, a.proc_tiles[band]);
    const double* in = a.in + (int64_t)band * a.in_band + (int64_t)chan * a.in_row;
    double* out = a.out ? a.out + (int64_t)band * a.out_band + (int64_t)chan * a.out_row : nullptr;

    double b0[S], b1[S], b2[S], a1[S], a2[S], z0[S], z1[S];
#pragma unroll
    for (int s = 0; s < S; ++s) {
        const double* c = a.coef + ((int64_t)band * S + s) * 5;
        b0[s] = c[0]; b1[s] = c[1]; b2[s] = c[2]; a1[s] = c[3]; a2[s] = c[4];
        z0[s] = 0.0; z1[s] = 0.0;
    }
    const int64_t mychunk = chunk0 + lane;
    const bool valid = !EDGE || mychunk < a.nchunks;
    const int64_t sidx = (((int64_t)band * a.C + chan) * a.nchunks + mychunk) * (2 * S);
    if (a.state_in && valid) {
        const int64_t rin = (((int64_t)band * a.C + chan) * a.sin_n + mychunk * a.sin_mul + a.sin_off) * (2 * S);
#pragma unroll
        for (int s = 0; s < S; ++s) { z0[s] = a.state_in[rin + 2 * s]; z1[s] = a.state_in[rin + 2 * s + 1]; }
    }

    // the copy this lane performs for each of its slots of a tile: row = slot*ROWS_PER_REQ + lane/SEGS
    const int seg = lane % FT_SEGS, row0 = lane / FT_SEGS;
    const int64_t lane_t0 = (chunk0 + row0) * a.L + seg * 2;      // sample index of slot 0, tile offset 0
    const int64_t slot_stride = (int64_t)FT_ROWS_PER_REQ * a.L;   // samples between consecutive slots
    double* const lane_dst = ring + row0 * FT_ROW + seg * 2;
    auto tile_off = [&](int i) -> int { return FT_TILE * (REV ? (ntiles - 1 - i) : i); };
    auto issue = [&](int i, int stage) {
        double* dst = lane_dst + stage * FT_TILE_DOUBLES;
        const int64_t t0 = lane_t0 + tile_off(i);
        const double* src = in + t0;
#pragma unroll
        for (int q = 0; q < FT_REQS; ++q) {
            if constexpr (!EDGE) {
                cp_async16(dst + q * FT_ROWS_PER_REQ * FT_ROW, src + q * slot_stride, 16);
            } else {
                const int64_t k = chunk0 + row0 + q * FT_ROWS_PER_REQ;
                const int64_t t = t0 + q * slot_stride;
                int bytes = 0;
                if (k < a.nchunks) {
                    const int64_t left = a.T - t;
                    bytes = left >= 2 ? 16 : (left == 1 ? 8 : 0);
                }
                cp_async16(dst + q * FT_ROWS_PER_REQ * FT_ROW, bytes ? (const void*)(in + t) : (const void*)in, bytes);
            }
        }
    };

#pragma unroll
    for (int s = 0; s < FT_STAGES - 1; ++s) {
        if (first + s < ntiles) issue(first + s, s);
        cp_async_commit();
    }
    int stage = 0;
    for (int i = first; i < ntiles; ++i) {
        {
            const int nxt = i + FT_STAGES - 1;
            int nstage = stage + FT_STAGES - 1;
            if (nstage >= FT_STAGES) nstage -= FT_STAGES;
            if (nxt < ntiles) issue(nxt, nstage);
            cp_async_commit();
        }
        cp_async_wait<FT_STAGES - 1>();
        __syncwarp();
        double* tile = ring + stage * FT_TILE_DOUBLES;
        double2* row = reinterpret_cast<double2*>(tile + lane * FT_ROW);
        double v[FT_TILE];
#pragma unroll
        for (int q = 0; q < FT_TILE / 2; ++q) {
            double2 d = row[q];
            v[2 * q] = d.x; v[2 * q + 1] = d.y;
        }
#pragma unroll
        for (int jj = 0; jj < FT_TILE; ++jj) {
            const int j = REV ? (FT_TILE - 1 - jj) : jj;
            double x = v[j];
#pragma unroll
            for (int s = 0; s < S; ++s) {
                if constexpr (STATE_ONLY || FUSED) {
                    // phase 1 only feeds the chunk-state scan, whose reassociation already differs from the
                    // sequential order by rounding: fused multiply-adds halve its FP64 instruction count.  The same
                    // holds for phase 3 of a multi-chunk run (its start states come from that scan); a single-chunk
                    // run keeps scipy's unfused order below and stays bit-identical to sosfilt
                    const double y = fma(b0[s], x, z0[s]);
                    z0[s] = fma(-a1[s], y, fma(b1[s], x, z1[s]));   // b1 x + z1 is off the z0 -> y -> z0 chain
                    z1[s] = fma(b2[s], x, -(a2[s] * y));
                    x = y;
                } else {
                    // scipy _sosfilt order, unfused
                    const double y = __dadd_rn(__dmul_rn(b0[s], x), z0[s]);
                    z0[s] = __dadd_rn(__dsub_rn(__dmul_rn(b1[s], x), __dmul_rn(a1[s], y)), z1[s]);
                    z1[s] = __dsub_rn(__dmul_rn(b2[s], x), __dmul_rn(a2[s], y));
                    x = y;
                }
            }
            v[j] = x;
        }
        if (!STATE_ONLY && out) {
#pragma unroll
            for (int q = 0; q < FT_TILE / 2; ++q) row[q] = make_double2(v[2 * q], v[2 * q + 1]);
            __syncwarp();
            const int64_t t0 = lane_t0 + tile_off(i);
            const double* srow = tile + row0 * FT_ROW + seg * 2;
#pragma unroll
            for (int q = 0; q < FT_REQS; ++q) {
                double2 d = *reinterpret_cast<const double2*>(srow + q * FT_ROWS_PER_REQ * FT_ROW);
                const int64_t t = t0 + q * slot_stride;
                if constexpr (MODE == 0) {
                    *reinterpret_cast<double2*>(out + t) = d;
                } else {
                    const int64_t k = chunk0 + row0 + q * FT_ROWS_PER_REQ;
                    if (k < a.nchunks && t < a.T) {
                        if (a.taper_len > 0) {
                            const int64_t g = a.g0 + t;
                            const int64_t dn = a.G - a.taper_len;
                            if (g < a.taper_len) d.x = __dmul_rn(d.x, a.taper_up[g]);
                            if (g + 1 < a.taper_len) d.y = __dmul_rn(d.y, a.taper_up[g + 1]);
                            if (g >= dn && g < a.G) d.x = __dmul_rn(d.x, a.taper_down[g - dn]);
                            if (g + 1 >= dn && g + 1 < a.G) d.y = __dmul_rn(d.y, a.taper_down[g + 1 - dn]);
                        }
                        if (t + 1 < a.T) *reinterpret_cast<double2*>(out + t) = d;
                        else out[t] = d.x;
                    }
                }
            }
        }
        __syncwarp();
        if (++stage == FT_STAGES) stage = 0;
    }
    cp_async_wait<0>();
    if (a.state_out && valid) {
#pragma unroll
        for (int s = 0; s < S; ++s) { a.state_out[sidx + 2 * s] = z0[s]; a.state_out[sidx + 2 * s + 1] = z1[s]; }
    }
}

template <int S, bool REV>
__global__ void __launch_bounds__(FT_WARPS * 32) biquad_chunk_kernel(FilterArgs a) {
    extern __shared__ __align__(16) double smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int band = blockIdx.x * FT_WARPS + warp;
    if (band >= a.B) return;   // warps are independent: only warp-level synchronisation below
    const int chan = blockIdx.z;
    const int64_t chunk0 = (int64_t)blockIdx.y * 32;
    double* ring = smem + warp * (FT_STAGES * FT_TILE_DOUBLES);
    // interior warp: its 32 chunks are whole (none is the record's ragged last chunk) and, when a taper is applied by
    // this launch, none of its samples lies in the first / last taper_len samples of the global record
    const int64_t s_lo = chunk0 * a.L, s_hi = (chunk0 + 32) * a.L;          // local sample range [s_lo, s_hi)
    const bool edge = s_hi > a.T;
    (void)s_lo;
    if (!a.out) {     // phase 1: chunk end states only
        if (edge) biquad_warp<S, REV, 2, true, true>(a, ring, band, chan, chunk0, lane);
        else biquad_warp<S, REV, 0, true, true>(a, ring, band, chan, chunk0, lane);
    } else if (a.nchunks > 1) {
        // whole chunks take the interior path even inside a taper zone: taper_rmw_kernel multiplies the ramps in
        // afterwards (the ragged last warp tapers its own samples)
        if (edge) biquad_warp<S, REV, 2, false, true>(a, ring, band, chan, chunk0, lane);
        else biquad_warp<S, REV, 0, false, true>(a, ring, band, chan, chunk0, lane);
    } else {
        biquad_warp<S, REV, 2, false, false>(a, ring, band, chan, chunk0, lane);
    }
}

// phase 2: state[k] (chunk end states from zero) -> state[k] (true chunk start states)
// Phase 1 of the forward pass when every band reads the same rows (the raw record): one block per (32 chunks, channel)
// moves each input tile ONCE through a block-wide cp.async ring and all bands take their chunk end states from it -
// warp w runs bands w, w+FS_WARPS, ... of the tile, NA of them at a time with the sample loop outermost, so a lane
// carries NA independent recurrences.  biquad_chunk_kernel would read the record once per band instead (its warps
// reach the same samples milliseconds apart, far beyond the residency of a line in L2).  Needs proc_tiles to be
// non-increasing over the band index (bands ordered by rising frequency), so that the active bands of a warp are a
// prefix of its list; the host checks that.  Same arithmetic per band as biquad_warp<.., STATE_ONLY>.
// FSW = warps per block (template parameter): 4, or 8 when more than 12 bands share a tile (then a warp carries at most
// three bands and takes them in ONE sweep per tile instead of two)
// bands [Q0, Q0 + NA) of the warp's list over one tile row
template <int S, int NBW, int Q0, int NA>
__device__ __forceinline__ void state_tile(const double* row, const double (*cf)[5], double (&z0)[NBW][S],
                                           double (&z1)[NBW][S]) {
    double b0[NA][S], b1[NA][S], b2[NA][S], a1[NA][S], a2[NA][S];
#pragma unroll
    for (int q = 0; q < NA; ++q)
#pragma unroll
        for (int s = 0; s < S; ++s) {
            const double* c = cf[(Q0 + q) * S + s];
            b0[q][s] = c[0]; b1[q][s] = c[1]; b2[q][s] = c[2]; a1[q][s] = c[3]; a2[q][s] = c[4];
        }
#pragma unroll 4
    for (int j = 0; j < FT_TILE; ++j) {
        const double xin = row[j];
#pragma unroll
        for (int q = 0; q < NA; ++q) {
            double x = xin;
#pragma unroll
            for (int s = 0; s < S; ++s) {
                const double y = fma(b0[q][s], x, z0[Q0 + q][s]);
                // b1 x + z1 does not wait for y: two dependent FMAs per sample on the z0 -> y -> z0 chain instead of three
                z0[Q0 + q][s] = fma(-a1[q][s], y, fma(b1[q][s], x, z1[Q0 + q][s]));
                z1[Q0 + q][s] = fma(b2[q][s], x, -(a2[q][s] * y));
                x = y;
            }
        }
    }
}

template <int S, int NBW, int FS_WARPS>
__global__ void __launch_bounds__(FS_WARPS * 32, 512 / (FS_WARPS * 32)) biquad_state_shared_kernel(FilterArgs a) {
    extern __shared__ __align__(16) double smem[];
    __shared__ double coef_sh[FS_WARPS][NBW * S][5];
    __shared__ int first_sh[FS_WARPS][NBW];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int chan = blockIdx.y;
    const int64_t chunk0 = (int64_t)blockIdx.x * 32;
    const int ntiles = a.L / FT_TILE;
    const double* in = a.in + (int64_t)chan * a.in_row;
    // this warp's bands, their coefficients and first tiles; the block starts at the earliest first tile of all bands
    const int bsplit = blockIdx.z, nsplit = a.nsplit;
    if (lane < NBW) {
        const int band = bsplit + nsplit * (warp + lane * FS_WARPS);
        first_sh[warp][lane] = band < a.B ? ntiles - min(ntiles, a.proc_tiles[band]) : ntiles;
    }
    for (int e = lane; e < NBW * S * 5; e += 32) {
        const int q = e / (S * 5), r = e - q * (S * 5);
        const int band = bsplit + nsplit * (warp + q * FS_WARPS);
        coef_sh[warp][q * S + r / 5][r % 5] = band < a.B ? a.coef[(int64_t)band * S * 5 + r] : 0.0;
    }
    __syncthreads();
    int first = ntiles;
#pragma unroll
    for (int w = 0; w < FS_WARPS; ++w) first = min(first, first_sh[w][0]);
    int myfirst[NBW];
#pragma unroll
    for (int q = 0; q < NBW; ++q) myfirst[q] = first_sh[warp][q];
    const bool edge = (chunk0 + 32) * a.L > a.T;

    // block-wide copy of tile i: 32 rows x FT_SEGS 16-byte pieces over FS_WARPS*32 threads
    auto issue = [&](int i, int stage) {
        double* tile = smem + stage * FT_TILE_DOUBLES;
#pragma unroll
        for (int q = 0; q < 32 * FT_SEGS / (FS_WARPS * 32); ++q) {
            const int id = tid + q * FS_WARPS * 32;
            const int r = id / FT_SEGS, seg = id - r * FT_SEGS;
            const int64_t t = (chunk0 + r) * a.L + (int64_t)FT_TILE * i + seg * 2;
            int bytes = 16;
            if (edge) {
                bytes = 0;
                if (chunk0 + r < a.nchunks) {
                    const int64_t left = a.T - t;
                    bytes = left >= 2 ? 16 : (left == 1 ? 8 : 0);
                }
            }
            cp_async16(tile + r * FT_ROW + seg * 2, bytes ? (const void*)(in + t) : (const void*)in, bytes);
        }
    };
    double z0[NBW][S], z1[NBW][S];
#pragma unroll
    for (int q = 0; q < NBW; ++q)
#pragma unroll
        for (int s_ = 0; s_ < S; ++s_) { z0[q][s_] = 0.0; z1[q][s_] = 0.0; }

#pragma unroll
    for (int st = 0; st < FT_STAGES - 1; ++st) {
        if (first + st < ntiles) issue(first + st, st);
        cp_async_commit();
    }
    int stage = 0;
    for (int i = first; i < ntiles; ++i) {
        {
            const int nxt = i + FT_STAGES - 1;
            int nstage = stage + FT_STAGES - 1;
            if (nstage >= FT_STAGES) nstage -= FT_STAGES;
            if (nxt < ntiles) issue(nxt, nstage);
            cp_async_commit();
        }
        cp_async_wait<FT_STAGES - 1>();
        __syncthreads();
        const double* row = smem + stage * FT_TILE_DOUBLES + lane * FT_ROW;
        // active bands of this warp at tile i: a prefix of its list (proc_tiles does not increase with the band index)
        int na = 0;
#pragma unroll
        for (int q = 0; q < NBW; ++q) na += (i >= myfirst[q]) ? 1 : 0;
        // NA recurrences side by side; the z arrays of the first NA bands are the leading rows
        // at most three recurrences side by side (their coefficients fill the registers), the rest in a second sweep
        auto sweep = [&](auto q0, auto na_) {
            constexpr int Q0 = decltype(q0)::value, NA = decltype(na_)::value;
            if constexpr (Q0 + NA <= NBW) state_tile<S, NBW, Q0, NA>(row, coef_sh[warp], z0, z1);
        };
        using std::integral_constant;
        switch (na) {
            case 1: sweep(integral_constant<int, 0>{}, integral_constant<int, 1>{}); break;
            case 2: sweep(integral_constant<int, 0>{}, integral_constant<int, 2>{}); break;
            case 3: sweep(integral_constant<int, 0>{}, integral_constant<int, 3>{}); break;
            case 4: sweep(integral_constant<int, 0>{}, integral_constant<int, 2>{});
                    sweep(integral_constant<int, 2>{}, integral_constant<int, 2>{}); break;
            case 5: sweep(integral_constant<int, 0>{}, integral_constant<int, 3>{});
                    sweep(integral_constant<int, 3>{}, integral_constant<int, 2>{}); break;
            case 6: sweep(integral_constant<int, 0>{}, integral_constant<int, 3>{});
                    sweep(integral_constant<int, 3>{}, integral_constant<int, 3>{}); break;
            default: break;
        }
        __syncthreads();          // every warp is done with this stage before the ring refills it
        if (++stage == FT_STAGES) stage = 0;
    }
    cp_async_wait<0>();
    const int64_t mychunk = chunk0 + lane;
    if (mychunk < a.nchunks) {
#pragma unroll
        for (int q = 0; q < NBW; ++q) {
            const int band = bsplit + nsplit * (warp + q * FS_WARPS);
            if (band < a.B) {
                const int64_t sidx = (((int64_t)band * a.C + chan) * a.nchunks + mychunk) * (2 * S);
#pragma unroll
                for (int s_ = 0; s_ < S; ++s_) { a.state_out[sidx + 2 * s_] = z0[q][s_]; a.state_out[sidx + 2 * s_ + 1] = z1[q][s_]; }
            }
        }
    }
}

// The Hann ramps of obspy's taper over the whole chunks of a multi-chunk run: y[t] *= ramp[g0 + t] for the local samples
// t < t_end that lie in the first / last taper_len samples of the GLOBAL record.  (Fused into the output pass these
// multiplies - two dependent table loads in front of every 16-byte store - cost the edge ranks of a time-sharded 8 x 24 h
// record 4.5 ms of a 12.8 ms kernel; as a read-modify-write over the zones alone they cost the zone's traffic.)  The same
// __dmul_rn as the fused path: identical bits.  grid = (blocks over the zone, C, bands); zone 0 = head, 1 = tail.
__global__ void __launch_bounds__(256) taper_rmw_kernel(double* y, int64_t row, int64_t band, int C, const double* up,
                                                        const double* down, int64_t taper_len, int64_t g0, int64_t G,
                                                        int64_t t_lo, int64_t t_hi, int zone) {
    const int b = blockIdx.z, c = blockIdx.y;
    (void)C;
    double* p = y + (int64_t)b * band + (int64_t)c * row;
    for (int64_t t = t_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < t_hi; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t g = g0 + t;
        p[t] = __dmul_rn(p[t], zone == 0 ? up[g] : down[g - (G - taper_len)]);
    }
}

template <int S2>
__global__ void carry_kernel(double* state, const double* ML, int B, int C, int nchunks, int reverse) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= B * C) return;
    const int band = id / C;
    double* st = state + (int64_t)id * nchunks * S2;
    // A^L of the band: in registers for up to 3 sections, otherwise read through L1 every step
    constexpr bool MREG = S2 <= 6;
    const double* Mg = ML + (int64_t)band * S2 * S2;
    double M[MREG ? S2 * S2 : 1], s[S2];
    if constexpr (MREG) {
#pragma unroll
        for (int i = 0; i < S2 * S2; ++i) M[i] = Mg[i];
    }
#pragma unroll
    for (int i = 0; i < S2; ++i) s[i] = 0.0;
    // the chunk end states e[k] do not depend on the recurrence: fetch PF of them ahead of the serial chain
    constexpr int PF = S2 <= 4 ? 8 : (S2 <= 8 ? 4 : 2);      // e[PF][S2] stays within 32 doubles
    for (int step0 = 0; step0 < nchunks; step0 += PF) {
        double e[PF][S2];
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int step = step0 + u;
            if (step < nchunks) {
                const double* p = st + (int64_t)(reverse ? nchunks - 1 - step : step) * S2;
#pragma unroll
                for (int i = 0; i < S2; ++i) e[u][i] = p[i];
            }
        }
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int step = step0 + u;
            if (step < nchunks) {
                double* p = st + (int64_t)(reverse ? nchunks - 1 - step : step) * S2;
                double t[S2];
#pragma unroll
                for (int i = 0; i < S2; ++i) {
                    p[i] = s[i];
                    double acc = e[u][i];
#pragma unroll
                    for (int j = 0; j < S2; ++j) acc = fma(MREG ? M[i * S2 + j] : __ldg(Mg + i * S2 + j), s[j], acc);
                    t[i] = acc;
                }
#pragma unroll
                for (int i = 0; i < S2; ++i) s[i] = t[i];
            }
        }
    }
}

typedef void (*state_fn)(FilterArgs);
// only the combinations pick_state() can select are instantiated: S * NBW <= 10 recurrence states per lane (more would
// not fit in 128 registers: those variants used to be compiled, spilling kilobytes, and never run)
template <int S, int NBW, int FSW>
static state_fn state_kernel_or_null() {
    if constexpr (S * NBW <= 10) return biquad_state_shared_kernel<S, NBW, FSW>;
    else return nullptr;
}
template <int S, int FSW>
static state_fn pick_state_nbw(int nbw) {
    switch (nbw) {
        case 1: return state_kernel_or_null<S, 1, FSW>();
        case 2: return state_kernel_or_null<S, 2, FSW>();
        case 3: return state_kernel_or_null<S, 3, FSW>();
        case 4: return state_kernel_or_null<S, 4, FSW>();
        case 5: return state_kernel_or_null<S, 5, FSW>();
        case 6: return state_kernel_or_null<S, 6, FSW>();
        default: return nullptr;
    }
}
template <int FSW>
static state_fn pick_state_w(int S, int nbw) {
    if (S * nbw > 10) return nullptr;      // more states than fit in 128 registers: the per-band kernel serves
    switch (S) {
        case 1: return pick_state_nbw<1, FSW>(nbw);
        case 2: return pick_state_nbw<2, FSW>(nbw);
        case 3: return pick_state_nbw<3, FSW>(nbw);
        case 4: return pick_state_nbw<4, FSW>(nbw);
        default: return nullptr;
    }
}
static state_fn pick_state(int S, int per_block, int fsw) {
    return fsw == 8 ? pick_state_w<8>(S, (per_block + 7) / 8) : pick_state_w<4>(S, (per_block + 3) / 4);
}

typedef void (*carry_fn)(double*, const double*, int, int, int, int);
static carry_fn pick_carry(int S) {
    switch (S) {
        case 1: return carry_kernel<2>;
        case 2: return carry_kernel<4>;
        case 3: return carry_kernel<6>;
        case 4: return carry_kernel<8>;
        case 5: return carry_kernel<10>;
        case 6: return carry_kernel<12>;
        case 7: return carry_kernel<14>;
        default: return carry_kernel<16>;
    }
}

typedef void (*chunk_fn)(FilterArgs);
template <int S>
static chunk_fn pick_dir(bool rev) { return rev ? biquad_chunk_kernel<S, true> : biquad_chunk_kernel<S, false>; }
static chunk_fn pick(int S, bool rev) {
    switch (S) {
        case 1: return pick_dir<1>(rev);
        case 2: return pick_dir<2>(rev);
        case 3: return pick_dir<3>(rev);
        case 4: return pick_dir<4>(rev);
        case 5: return pick_dir<5>(rev);
        case 6: return pick_dir<6>(rev);
        case 7: return pick_dir<7>(rev);
        case 8: return pick_dir<8>(rev);
    }
    return nullptr;
}

// one zero-input step of the DF2T cascade as a (2S x 2S) matrix, then its L-th power
static void transition_power(const double* sos, int S, int64_t L, std::vector<double>& out) {
    const int n = 2 * S;
    std::vector<long double> M(n * n, 0.0L), R(n * n, 0.0L), Tm(n * n);
    for (int col = 0; col < n; ++col) {
        std::vector<long double> z(n, 0.0L);
        z[col] = 1.0L;
        long double x = 0.0L;
        for (int s = 0; s < S; ++s) {
            const long double b0 = sos[s * 6 + 0], b1 = sos[s * 6 + 1], b2 = sos[s * 6 + 2];
            const long double a1 = sos[s * 6 + 4], a2 = sos[s * 6 + 5];
            const long double y = b0 * x + z[2 * s];
            const long double n0 = b1 * x - a1 * y + z[2 * s + 1];
            const long double n1 = b2 * x - a2 * y;
            z[2 * s] = n0; z[2 * s + 1] = n1;
            x = y;
        }
        for (int r = 0; r < n; ++r) M[r * n + col] = z[r];
    }
    for (int i = 0; i < n; ++i) R[i * n + i] = 1.0L;
    auto mul = [&](std::vector<long double>& A, const std::vector<long double>& Bm) {
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                long double acc = 0.0L;
                for (int k = 0; k < n; ++k) acc += A[i * n + k] * Bm[k * n + j];
                Tm[i * n + j] = acc;
            }
        A = Tm;
    };
    int64_t e = L;
    while (e > 0) {
        if (e & 1) mul(R, M);
        e >>= 1;
        if (e) { std::vector<long double> M2 = M; mul(M, M2); }
    }
    out.resize(n * n);
    for (int i = 0; i < n * n; ++i) out[i] = (double)R[i];
}

static double max_pole_radius(const double* sos, int S) {
    double r = 0.0;
    for (int s = 0; s < S; ++s) {
        const double a1 = sos[s * 6 + 4], a2 = sos[s * 6 + 5];
        const double disc = a1 * a1 - 4.0 * a2;
        double m;
        if (disc < 0) m = sqrt(fabs(a2));
        else m = std::max(fabs((-a1 + sqrt(disc)) * 0.5), fabs((-a1 - sqrt(disc)) * 0.5));
        r = std::max(r, m);
    }
    return r;
}

int filter_impl(nbls_ctx* ctx, const nbls_filter_desc* d, const double* x, int C, int64_t T, int64_t ldx,
                double* y, int64_t ldy, int64_t ystride_band) {
    NBLS_REQUIRE(d && x && y, "nbls_filter: null argument");
    const int B = d->nbands, S = d->nsections;
    NBLS_REQUIRE(B >= 1 && B <= NBLS_MAX_BANDS, "nbls_filter: nbands %d out of range", B);
    NBLS_REQUIRE(S >= 1 && S <= NBLS_MAX_SECTIONS, "nbls_filter: nsections %d out of range [1,%d]", S,
                 NBLS_MAX_SECTIONS);
    NBLS_REQUIRE(C >= 1 && C <= 65535 && T >= 1, "nbls_filter: bad shape C=%d T=%lld", C, (long long)T);
    NBLS_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0 && (ldx & 1) == 0 && (ldy & 1) == 0 &&
                     (ystride_band & 1) == 0,
                 "nbls_filter: rows must be 16-byte aligned (even strides, aligned base)");
    NBLS_REQUIRE(ldx >= T && ldy >= T, "nbls_filter: row stride shorter than npts");
    for (int i = 0; i < B * S; ++i)
        NBLS_REQUIRE(d->sos[i * 6 + 3] == 1.0, "nbls_filter: sos[:,3] must be 1 (scipy sosfilt contract)");
    NBLS_REQUIRE(d->taper_len >= 0 && (d->taper_len == 0 || (d->taper_up && d->taper_down)),
                 "nbls_filter: taper arrays missing");

    // chunking
    int64_t L = d->chunk_len;
    // The forward state pass (phase 1) can run on sub-chunks of L / sub samples (sub x the lanes, recurrence chains sub x
    // shorter; the carry then walks sub x the states and phase 3 picks every sub-th one).  Measured on c3/24 h (B200):
    // sub = 1 / 2 / 4 / 8 -> 2.96 / 3.07 / 3.61 / 3.71 ms: the pass is bound by its FP64 work, which grows with sub (every
    // sub-chunk reruns its K_b tail), not by the latency of the chains - the default stays 1; NBLS_OPT_FILTER_P1_SUB selects.
    int sub = (int)ctx->opt[NBLS_OPT_FILTER_P1_SUB];
    if (sub != 1 && sub != 2 && sub != 4 && sub != 8) sub = 1;
    if (d->exact_scan) sub = 1;
    if (L <= 0) {
        const double lanes = (double)ctx->opt[NBLS_OPT_FILTER_LANES];
        L = round_up((int64_t)((double)T * B * C / lanes), FT_TILE * sub);
        L = std::max<int64_t>(256, std::min<int64_t>(65536, L));
    }
    while (sub > 1 && (L % (FT_TILE * sub) != 0 || L / sub < 4 * FT_TILE)) sub >>= 1;
    NBLS_REQUIRE(L % FT_TILE == 0 && L >= FT_TILE, "nbls_filter: chunk_len must be a positive multiple of %d", FT_TILE);
    L = std::min<int64_t>(L, round_up(T, FT_TILE));
    const int64_t nchunks64 = (T + L - 1) / L;
    NBLS_REQUIRE(nchunks64 <= (1 << 26), "nbls_filter: too many chunks");
    const int nchunks = (int)nchunks64;
    const int ntiles = (int)(L / FT_TILE);
    const int S2 = 2 * S;
    if (nchunks <= 1) sub = 1;
    const int64_t L1 = L / sub;                                  // sub-chunk length of the forward state pass
    const int nchunks1 = (int)((T + L1 - 1) / L1), ntiles1 = (int)(L1 / FT_TILE);

    // host-side tables: coefficients, A^L, phase-1 lengths
    const size_t n_coef = (size_t)B * S * 5, n_ml = (size_t)B * S2 * S2;
    std::vector<double> hbuf(n_coef + 2 * n_ml);                 // coefficients, A^L (backward carry), A^L1 (forward carry)
    std::vector<int> hproc(2 * B);                               // phase-1 tiles per chunk: backward [B], forward [B]
    for (int b = 0; b < B; ++b) {
        const double* sb = d->sos + (size_t)b * S * 6;
        for (int s = 0; s < S; ++s) {
            double* c = &hbuf[((size_t)b * S + s) * 5];
            c[0] = sb[s * 6 + 0]; c[1] = sb[s * 6 + 1]; c[2] = sb[s * 6 + 2];
            c[3] = sb[s * 6 + 4]; c[4] = sb[s * 6 + 5];
        }
        std::vector<double> ml;
        transition_power(sb, S, L, ml);
        memcpy(&hbuf[n_coef + (size_t)b * S2 * S2], ml.data(), sizeof(double) * S2 * S2);
        transition_power(sb, S, L1, ml);
        memcpy(&hbuf[n_coef + n_ml + (size_t)b * S2 * S2], ml.data(), sizeof(double) * S2 * S2);
        int pt = ntiles, pt1 = ntiles1;
        if (!d->exact_scan) {
            const double r = max_pole_radius(sb, S);
            if (r > 0.0 && r < 1.0) {
                // K from the spectral radius, then grown until the transition matrix itself has decayed: the cascade of a
                // narrow band-pass is strongly non-normal (its poles cluster), so ||A^K|| can exceed r^K by orders of magnitude
                double K = ceil(64.0 * M_LN2 / -log(r));
                const double target = ldexp(1.0, -64);
                for (int grow = 0; grow < 64 && K < (double)L; ++grow) {
                    std::vector<double> ak;
                    transition_power(sb, S, (int64_t)K, ak);
                    double nrm = 0.0;
                    for (int i = 0; i < S2; ++i) {
                        double row = 0.0;
                        for (int jn = 0; jn < S2; ++jn) row += fabs(ak[i * S2 + jn]);
                        nrm = std::max(nrm, row);
                    }
                    if (nrm <= target) break;
                    K = ceil(K + std::max(32.0, log(nrm / target) / -log(r)));
                }
                if (K < (double)L) pt = (int)std::min<double>(ntiles, ceil(K / FT_TILE));
                if (K < (double)L1) pt1 = (int)std::min<double>(ntiles1, ceil(K / FT_TILE));
            } else if (r == 0.0) {
                pt = std::min(ntiles, 2 * S / FT_TILE + 1);   // FIR: state forgets after 2 samples/section
                pt1 = std::min(ntiles1, 2 * S / FT_TILE + 1);
            }
        }
        hproc[b] = pt;
        hproc[B + b] = pt1;
    }
    const size_t coef_bytes = hbuf.size() * sizeof(double) + round_up(2 * B * sizeof(int), 16);
    NBLS_CHECK(ensure(ctx, ctx->coef, coef_bytes));
    double* dcoef = (double*)ctx->coef.p;
    double* dml = dcoef + n_coef;
    double* dml1 = dml + n_ml;
    int* dproc = (int*)(dml1 + n_ml);
    int* dproc1 = dproc + B;
    // The device copies of the tables are kept between calls: a caller that repeats a job (a queue of records with one
    // band plan) pays the two uploads and their synchronisations - a few hundred microseconds of idle GPU - only once.
    // (compared segment by segment against the kept host image: no allocation and one memcmp pass - the taper ramps of a
    // time-sharded multi-day record are megabytes)
    const size_t k_coef = hbuf.size() * sizeof(double), k_proc = hproc.size() * sizeof(int);
    const size_t k_tap = (size_t)d->taper_len * sizeof(double);
    const std::vector<unsigned char>& fk = ctx->filter_key;
    bool cached = fk.size() == k_coef + k_proc + 2 * k_tap && ctx->filter_key_coef == ctx->coef.p &&
                  (d->taper_len == 0 || ctx->filter_key_taper == ctx->taper.p);
    cached = cached && memcmp(fk.data(), hbuf.data(), k_coef) == 0 && memcmp(fk.data() + k_coef, hproc.data(), k_proc) == 0;
    if (cached && d->taper_len > 0)
        cached = memcmp(fk.data() + k_coef + k_proc, d->taper_up, k_tap) == 0 &&
                 memcmp(fk.data() + k_coef + k_proc + k_tap, d->taper_down, k_tap) == 0;
    const double *dup = nullptr, *ddown = nullptr;
    if (d->taper_len > 0) NBLS_CHECK(ensure(ctx, ctx->taper, 2 * d->taper_len * sizeof(double)));
    if (!cached || (d->taper_len > 0 && ctx->filter_key_taper != ctx->taper.p)) {
        NBLS_CUDA(cudaMemcpyAsync(dcoef, hbuf.data(), hbuf.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        NBLS_CUDA(cudaMemcpyAsync(dproc, hproc.data(), 2 * B * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        if (d->taper_len > 0) {
            double* tp = (double*)ctx->taper.p;
            NBLS_CUDA(cudaMemcpyAsync(tp, d->taper_up, d->taper_len * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
            NBLS_CUDA(cudaMemcpyAsync(tp + d->taper_len, d->taper_down, d->taper_len * sizeof(double),
                                      cudaMemcpyHostToDevice, ctx->stream));
        }
        // pageable sources: wait until the copies have left the host buffers
        NBLS_CUDA(cudaStreamSynchronize(ctx->stream));
        std::vector<unsigned char>& key = ctx->filter_key;
        key.resize(k_coef + k_proc + 2 * k_tap);
        memcpy(key.data(), hbuf.data(), k_coef);
        memcpy(key.data() + k_coef, hproc.data(), k_proc);
        if (d->taper_len > 0) {
            memcpy(key.data() + k_coef + k_proc, d->taper_up, k_tap);
            memcpy(key.data() + k_coef + k_proc + k_tap, d->taper_down, k_tap);
        }
        ctx->filter_key_coef = ctx->coef.p;
        ctx->filter_key_taper = d->taper_len > 0 ? ctx->taper.p : nullptr;
    }
    if (d->taper_len > 0) {
        dup = (const double*)ctx->taper.p;
        ddown = dup + d->taper_len;
    }
    double* dstate = nullptr;
    if (nchunks > 1) {
        NBLS_CHECK(ensure(ctx, ctx->state, (size_t)B * C * nchunks1 * S2 * sizeof(double)));
        dstate = (double*)ctx->state.p;
    }

    const size_t smem = (size_t)FT_WARPS * FT_STAGES * FT_TILE_DOUBLES * sizeof(double);
    NBLS_REQUIRE((nchunks1 + 31) / 32 <= 65535, "nbls_filter: chunk grid too large (raise chunk_len)");
    for (int rev = 0; rev < 2; ++rev) {
        chunk_fn fn = pick(S, rev != 0);
        NBLS_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    bool mono = true;
    for (int b = 1; b < B; ++b) mono = mono && hproc[B + b] <= hproc[B + b - 1];

    // One pass (forward or backward) over the bands [b0, b0 + nb) on stream `st`: state pass, carry, output pass.
    auto run_pass = [&](bool rev, const double* in, int64_t in_row, int64_t in_band, bool taper, int b0, int nb,
                        cudaStream_t st) -> int {
        chunk_fn fn = pick(S, rev);
        const dim3 grid((nb + FT_WARPS - 1) / FT_WARPS, (nchunks + 31) / 32, C);
        // the state pass of the forward direction may run on sub-chunks (see above), the backward one on whole chunks
        const int psub = rev ? 1 : sub;
        const int pn = rev ? nchunks : nchunks1;
        double* dstate_g = dstate ? dstate + (size_t)b0 * C * nchunks1 * S2 : nullptr;     // every band owns nchunks1 slots
        FilterArgs a;
        memset(&a, 0, sizeof(a));
        a.in = in + (int64_t)b0 * in_band; a.in_row = in_row; a.in_band = in_band;
        a.coef = dcoef + (size_t)b0 * S * 5; a.B = nb; a.C = C; a.T = T;
        if (nchunks > 1) {
            a.L = (int)(rev ? L : L1); a.nchunks = pn;
            a.out = nullptr; a.proc_tiles = (rev ? dproc : dproc1) + b0; a.state_in = nullptr; a.state_out = dstate_g;
            const dim3 grid1(grid.x, (pn + 31) / 32, C);
            {
                KernelGroupScope ks(ctx, rev ? KG_FILTER_P1B : KG_FILTER_P1F, 0.0);   // state passes: no algorithmic bytes
                // shared-read phase 1 of the forward pass (every band reads the same record: one read of each tile for all
                // bands of the group); needs the bands' decay lengths to fall with the band index.  The bands can be dealt
                // to `nsplit` blocks per (32 chunks, channel).
                state_fn shared_p1 = nullptr;
                int p1_warps = 4;
                const int nsplit = (int)std::max<int64_t>(1, ctx->opt[NBLS_OPT_FILTER_SPLIT]);
                if (!rev && in_band == 0 && mono && ctx->opt[NBLS_OPT_FILTER_SHARED_P1]) {
                    const int per_block = (nb + nsplit - 1) / nsplit;
                    // warps per block: 8 when 4 would leave a warp with more than three bands (two sweeps per tile), else 4
                    p1_warps = (int)ctx->opt[NBLS_OPT_FILTER_P1_WARPS];
                    if (p1_warps != 4 && p1_warps != 8) p1_warps = ((per_block + 3) / 4 > 3) ? 8 : 4;
                    shared_p1 = pick_state(S, per_block, p1_warps);
                    if (!shared_p1 && p1_warps == 4) { p1_warps = 8; shared_p1 = pick_state(S, per_block, 8); }
                }
                if (shared_p1) {
                    const size_t smem1 = (size_t)FT_STAGES * FT_TILE_DOUBLES * sizeof(double);
                    a.nsplit = nsplit;
                    shared_p1<<<dim3(grid1.y, C, nsplit), p1_warps * 32, smem1, st>>>(a);
                } else {
                    fn<<<grid1, FT_WARPS * 32, smem, st>>>(a);
                }
            }
            NBLS_CUDA(cudaGetLastError());
            {
                KernelGroupScope ks(ctx, KG_FILTER_CARRY, 0.0);
                pick_carry(S)<<<(nb * C + 31) / 32, 32, 0, st>>>(dstate_g, (rev ? dml : dml1) + (size_t)b0 * S2 * S2, nb, C, pn,
                                                                 rev ? 1 : 0);
            }
            NBLS_CUDA(cudaGetLastError());
            ctx->launches += 2;
        }
        a.L = (int)L; a.nchunks = nchunks;
        a.out = y + (int64_t)b0 * ystride_band; a.out_row = ldy; a.out_band = ystride_band;
        a.proc_tiles = nullptr; a.state_in = nchunks > 1 ? dstate_g : nullptr; a.state_out = nullptr;
        a.sin_n = pn; a.sin_mul = psub; a.sin_off = 0;
        if (taper) {
            a.taper_up = dup; a.taper_down = ddown; a.taper_len = d->taper_len; a.g0 = d->global_t0;
            a.G = d->global_npts > 0 ? d->global_npts : T;
        }
        {
            // algorithmic bytes: the raw record is read once for all bands (in_band == 0), a filtered band once; one write
            const double rd = in_band == 0 ? 8.0 * C * (double)T * nb / B : 8.0 * nb * C * (double)T;
            KernelGroupScope ks(ctx, rev ? KG_FILTER_P3B : KG_FILTER_P3F, rd + 8.0 * nb * C * (double)T);
            fn<<<grid, FT_WARPS * 32, smem, st>>>(a);
        }
        NBLS_CUDA(cudaGetLastError());
        ctx->launches += 1;
        if (taper && d->taper_len > 0 && nchunks > 1) {
            // whole chunks were written untapered: local samples [0, t_whole) of the two global taper zones
            const int64_t t_whole = std::min<int64_t>(T, (T / (32 * L)) * (32 * L));      // below the ragged last warp
            const int64_t G = a.G, g0 = a.g0, wl = d->taper_len;
            const int64_t h_lo = 0, h_hi = std::min<int64_t>(t_whole, wl - g0);                         // g < wl
            const int64_t t_lo2 = std::max<int64_t>(0, G - wl - g0), t_hi2 = std::min<int64_t>(t_whole, G - g0);   // g >= G - wl
            for (int zone = 0; zone < 2; ++zone) {
                const int64_t lo = zone == 0 ? h_lo : t_lo2, hi = zone == 0 ? h_hi : t_hi2;
                if (hi <= lo) continue;
                KernelGroupScope ks(ctx, rev ? KG_FILTER_P3B : KG_FILTER_P3F, 0.0);
                const unsigned gx = (unsigned)std::min<int64_t>(1024, (hi - lo + 255) / 256);
                taper_rmw_kernel<<<dim3(gx, (unsigned)C, (unsigned)nb), 256, 0, st>>>(a.out, ldy, ystride_band, C, dup, ddown, wl, g0, G,
                                                                           lo, hi, zone);
                NBLS_CUDA(cudaGetLastError());
                ctx->launches += 1;
            }
        }
        return NBLS_OK;
    };
    auto run_group = [&](int b0, int nb, cudaStream_t st) -> int {
        NBLS_CHECK(run_pass(false, x, ldx, 0, !d->zerophase, b0, nb, st));
        if (d->zerophase) NBLS_CHECK(run_pass(true, y, ldy, ystride_band, true, b0, nb, st));
        return NBLS_OK;
    };

    // Two band groups on two streams: the state passes are bound by FP64 work at low occupancy (the forward one runs 225
    // blocks on 148 SMs), the output passes by HBM bandwidth - side by side the low group's state passes hide under the
    // high group's output passes.  The low group holds the bands whose state passes are long (it is enqueued first, so
    // its few blocks are resident before the other stream fills the SMs).  Not in profile mode: there every group is
    // timed on its own.
    int b_lo = 0;
    if (d->zerophase && nchunks > 1 && B >= 8 && ctx->opt[NBLS_OPT_FILTER_OVERLAP] && !ctx->opt[NBLS_OPT_PROFILE]) {
        int64_t tot = 0, acc = 0;
        for (int b = 0; b < B; ++b) tot += hproc[b];
        // low group = the bands that hold ~40 % of the state-pass work (option values >= 10 give the percentage), at
        // least 2 and at most B - 2 bands.  Measured on c3/24 h (filter stage, ms): one stream 28.04; 12-30 % 28.1-27.8;
        // 40 % 26.3-26.4; 55 % 26.8; 70 % 26.9; 85-95 % 27.5
        const int64_t pct = ctx->opt[NBLS_OPT_FILTER_OVERLAP] >= 10 ? std::min<int64_t>(95, ctx->opt[NBLS_OPT_FILTER_OVERLAP]) : 40;
        for (b_lo = 0; b_lo < B && acc * 100 < tot * pct; ++b_lo) acc += hproc[b_lo];
        b_lo = std::max(2, std::min(B - 2, b_lo));
    }
    if (b_lo == 0) return run_group(0, B, ctx->stream);
    if (!ctx->stream2) {
        NBLS_CUDA(cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
        NBLS_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
        NBLS_CUDA(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    }
    NBLS_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream));
    NBLS_CUDA(cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0));
    NBLS_CHECK(run_group(0, b_lo, ctx->stream));
    NBLS_CHECK(run_group(b_lo, B - b_lo, ctx->stream2));
    NBLS_CUDA(cudaEventRecord(ctx->ev_join, ctx->stream2));
    NBLS_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    return NBLS_OK;
}

}  // namespace nbls
