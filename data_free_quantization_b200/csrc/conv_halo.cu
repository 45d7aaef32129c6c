// Kernel (e), static-range form: dense 3 x 3 convolution, stride 1, dilation 1, padding 1 (DeepLab's decoder:
// modeling/segmentation/decoder.py:22-30 under QuantNConv2d.forward, utils/quantize.py:250-256) WITHOUT an im2col gather.
//
// A tile is 16 x 8 output pixels.  Its 18 x 10 input halo -- all channels, as centred s8 codes -- is loaded ONCE into
// shared memory as [16-byte channel chunk][halo row][halo column][16 B] by six loader warps (a thread per halo pixel:
// 128-bit loads of its channel chunks, conflict-free 128-bit stores; a 4-D tensor-map box {16 B, 10, 18, 1} per chunk did
// the same transposition by TMA but took ~50 k cycles per tile -- 3600 sixteen-byte rows).  Pixels outside the image are
// zero bytes (the zero padding, corrected in the epilogue as in pw_i8.cu).
// In that layout the A operand of tap (i, j) is the SAME buffer read through another descriptor: eight consecutive
// columns of a halo row are a contiguous 8 x 16 B core matrix, the eight-row groups of the 128 GEMM rows are the tile's
// rows (stride = one halo row = 160 B, SBO), the K direction steps one chunk (2944 B, LBO), and the tap only moves the
// start address by (i * 10 + j) * 16 B.  So the nine taps cost NO activation traffic and no producer work: the register
// gather of pw_i8_kernel<kConv, kQ> re-read every input pixel nine times from L2 and spent ~110 warp instructions per
// 64-byte slab and warp on it (profiles/r2_conv_q8_ncu.md).
//
//   warp 0   (one lane) the weight slabs (bulk copies, ring)
//   warp 1   (one lane) tcgen05.mma.kind::i8, M = 128, N = channels of the pass, two TMEM accumulators
//   warps 2-17 epilogue: SUM_k a from the per-pixel code sums, tcgen05.ld, bias / ReLU, exact requantisation to the consumer
//            grid(s), 16 channels of a pixel per 128-bit store
//   warps 18-23 halo loaders, one tile ahead (two halo buffers); they also form the per-pixel code sums (dp4a)
// Weights: the packed layout of dfq_pw_i8_pack_cap_host over a reduction ordered [tap][chunk], chunks per tap padded to a
// multiple of 4 (a 64-byte weight slab never straddles two taps).
#include <cstdlib>
#include "common.cuh"
#include "q8_pack.cuh"
#include "q8_ptx.cuh"

namespace dfq {
namespace {

constexpr int kHTY = 16, kHTX = 8;                  // tile: 16 rows x 8 columns = 128 pixels (one 8-row group per tile row)
constexpr int kHY = kHTY + 2, kHX = kHTX + 2;       // halo
constexpr int kHBox = kHY * kHX * 16;               // bytes of one channel chunk of the halo: 2880 (one TMA box)
constexpr int kHChunk = (kHBox + 127) / 128 * 128;  // its stride in shared memory: TMA destinations are 128-byte aligned
constexpr int kHEpiWarps = 16;                      // the requantising epilogue is the long pole: 16 warps share a tile
constexpr int kHLoaders = 6;
constexpr int kHThreads = 32 * (2 + kHEpiWarps + kHLoaders);   // warps: 0 weights, 1 MMA, 2-17 epilogue, 18-23 halo loaders
constexpr int kHSlab = 64;                          // K bytes per weight slab (pw_i8 packing): 4 chunks, 2 MMAs
constexpr int kHMaxStages = 8;

struct HaloArgs {
    const int8_t *x; const int8_t *wpk; const float *wr; const float *wrt; const float *bias; const RangeSlot *in_slot;
    const RangeSlot *out_slot[2]; int8_t *y[2]; int cp_out;
    float *yf; unsigned *stats;         // kF32 (dynamic-range plan): fp32 NCHW output + encoded per-sample extrema
    int N, C, H, W, O, relu, n_pass, o_per_pass, npass, slabs, cpt, tiles_x, tiles_y, tmem_cols, stages;
    int64_t tiles;
    float s_w, c_w;
};

__device__ __forceinline__ bool elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}

// Launched as clusters of TWO CTAs (neighbouring SMs) that work on different tiles of the same pass and SHARE the weight
// stream: each CTA copies half of every slab and multicasts it into both shared memories, and a ring stage is refilled
// only when the MMAs of both CTAs have retired it (tcgen05.commit multicast onto both `bempty` barriers).  At one CTA per
// SM the weights were 720 KB per tile against 58 KB of activations and the kernel sat on the L2 -> SM bandwidth this
// GPU delivered (8.9 TB/s, tensor pipe 51 %).  Both CTAs of a pair run the same number of tile iterations (a tile index
// past the end is a ghost: loads and stores are skipped, the barriers are still walked).
template <bool kF32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kHThreads, 1)
conv3x3_halo_kernel(HaloArgs a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kHMaxStages + 8];   // bfull | bempty | hfull[2] hempty[2] tfull[2] tempty[2]
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float chan_c[256];
    __shared__ int S_s[2][kHY * kHX];                             // per halo pixel: sum over channels of the codes
    __shared__ uint32_t a_off[128];                               // per weight slab: where its first chunk's A operand starts
                                                                  // inside a halo buffer, in 16-byte units (tap shift + chunk)

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    const uint32_t halo_bytes = (uint32_t)a.cpt * kHChunk;
    const uint32_t bslab = (uint32_t)a.npass * kHSlab;
    const uint32_t sH_u = smem_u32(smem), sB_u = sH_u + 2 * halo_bytes;
    float *tap_c = reinterpret_cast<float *>(smem + 2 * halo_bytes + (size_t)a.stages * 2 * bslab);
    const uint32_t bfull0 = smem_u32(&bars[0]), bempty0 = bfull0 + 8 * kHMaxStages, hfull0 = bempty0 + 8 * kHMaxStages,
                   hempty0 = hfull0 + 16, tfull0 = hempty0 + 16, tempty0 = tfull0 + 16;
    const uint32_t crank = cluster_ctarank();                    // 0 / 1 inside the pair
    const int pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;
    const int pass = pair % a.n_pass;
    const int64_t tstep = (int64_t)(pairs / a.n_pass) * 2;
    const int64_t tile0 = (int64_t)(pair / a.n_pass) * 2 + crank;
    const int64_t tile_end = (a.tiles + tstep - 1) / tstep * tstep;   // every CTA walks the same number of iterations
    const int o0 = pass * a.o_per_pass, n_real = min(a.o_per_pass, a.O - o0);
    const int per_img = a.tiles_x * a.tiles_y;

    if (tid == 0) {
        for (int i = 0; i < a.stages; ++i) { mbar_init(bfull0 + 8 * i, 1); mbar_init(bempty0 + 8 * i, 2); }   // bempty: both CTAs' commits
        for (int i = 0; i < 2; ++i) {
            mbar_init(hfull0 + 8 * i, kHLoaders); mbar_init(hempty0 + 8 * i, 1 + kHEpiWarps);   // hempty: one commit + the epilogue warps
            mbar_init(tfull0 + 8 * i, 1); mbar_init(tempty0 + 8 * i, kHEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const RangeSlot in = *a.in_slot;
    const float sx = in.scale;
    const float cx = __fmaf_rn(128.0f, sx, in.mn);                // x_hat = sx * a + cx with a = code - 128
    if (tid < 256) {
        float c = 0.0f;
        if (tid < n_real) {
            const int o = o0 + tid;
            c = __fmaf_rn(cx, __ldg(a.wr + o), __fmul_rn(__fmul_rn((float)(9 * a.C), a.c_w), cx));
            if (a.bias) c = __fadd_rn(c, __ldg(a.bias + o));
        }
        chan_c[tid] = c;
    }
    for (int i = tid; i < a.slabs; i += kHThreads) {
        const int q = 4 * i, tap = q / a.cpt, cc = q - tap * a.cpt;
        a_off[i] = (uint32_t)(cc * (kHChunk >> 4) + (tap / 3) * kHX + tap % 3);
    }
    {   // what one out-of-image tap contributed through c_x, per output channel of this pass
        const float ccw = __fmul_rn(__fmul_rn((float)a.C, a.c_w), cx);
        for (int i = tid; i < 9 * a.npass; i += kHThreads) {
            const int t = i / a.npass, ol = i - t * a.npass;
            tap_c[i] = ol < n_real ? __fmaf_rn(cx, __ldg(a.wrt + (int64_t)t * a.O + o0 + ol), ccw) : 0.0f;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    cluster_sync();                                              // the peer's barriers exist before anything is sent to them
    const uint32_t tmem = tmem_base_s;

    auto tile_origin = [&](int64_t tile, int &n, int &oh0, int &ow0) {
        n = (int)(tile / per_img);
        const int r = (int)(tile - (int64_t)n * per_img);
        oh0 = (r / a.tiles_x) * kHTY; ow0 = (r % a.tiles_x) * kHTX;
    };

    if (warp == 0) {
        // ===== producer
        if (lane == 0) {
            const int8_t *wsrc = a.wpk + (int64_t)pass * a.slabs * bslab;
            int stage = 0; uint32_t phase = 0;
            for (int64_t tile = tile0; tile < tile_end; tile += tstep) {
                for (int s = 0; s < a.slabs; s += 2) {               // a ring stage = two consecutive 64-byte slabs
                    const uint32_t bytes = (uint32_t)min(2, a.slabs - s) * bslab;
                    mbar_wait(bempty0 + 8 * stage, phase ^ 1);
                    mbar_expect_tx(bfull0 + 8 * stage, bytes);                  // this CTA's half + the peer's half
                    bulk_g2s_mc(sB_u + stage * 2 * bslab + crank * (bytes / 2),
                                wsrc + (int64_t)s * bslab + crank * (bytes / 2), bytes / 2, bfull0 + 8 * stage, (uint16_t)3);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer.  The whole warp walks the loop (all its values are warp-uniform, so the compiler keeps the
        // descriptors in uniform registers) and one elected lane issues: with the loop inside `if (lane == 0)` every
        // tcgen05 instruction was wrapped in an ELECT / 7 x R2UR loop.
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.npass >> 3) << 17) |
                               ((uint32_t)(128 >> 4) << 24);
        int stage = 0; uint32_t phase = 0;
        int acc = 0; uint32_t aphase = 0;
        int hb = 0; uint32_t hphase = 0;
        const uint64_t bd0 = desc_plain(sB_u, a.npass * 16, 128);
        for (int64_t tile = tile0; tile < tile_end; tile += tstep) {
            mbar_wait(tempty0 + 8 * acc, aphase ^ 1);
            mbar_wait(hfull0 + 8 * hb, hphase);
            tc_fence_after();
            const uint32_t d = tmem + (uint32_t)(acc * a.npass);
            const uint64_t ad0 = desc_plain(sH_u + hb * halo_bytes, kHChunk, kHX * 16);
            for (int s = 0; s < a.slabs; s += 2) {
                mbar_wait(bfull0 + 8 * stage, phase);
                tc_fence_after();
                const uint64_t bd = bd0 + (uint64_t)(stage * (2 * bslab >> 4));
                const uint64_t ada = ad0 + a_off[s];
                const bool two = s + 1 < a.slabs;
                const uint64_t adb = ad0 + a_off[two ? s + 1 : s];
                if (elect_one()) {
                    umma_i8(d, ada, bd, idesc, s != 0);
                    umma_i8(d, ada + 2 * (kHChunk >> 4), bd + (uint64_t)(2 * a.npass), idesc, 1);
                    if (two) {
                        umma_i8(d, adb, bd + (uint64_t)(bslab >> 4), idesc, 1);
                        umma_i8(d, adb + 2 * (kHChunk >> 4), bd + (uint64_t)((bslab >> 4) + 2 * a.npass), idesc, 1);
                    }
                    umma_commit_mc(bempty0 + 8 * stage, (uint16_t)3);
                }
                __syncwarp();
                if (++stage == a.stages) { stage = 0; phase ^= 1; }
            }
            if (elect_one()) {
                umma_commit(tfull0 + 8 * acc);
                umma_commit(hempty0 + 8 * hb);
            }
            __syncwarp();
            if (++acc == 2) { acc = 0; aphase ^= 1; }
            hb ^= 1; if (hb == 0) hphase ^= 1;
        }
    } else if (warp < 2 + kHEpiWarps) {
        // ===== epilogue warps: TMEM lanes 32 * (w % 4) .., the four warps of a lane group take every fourth 16-column chunk
        const int lg = warp & 3, quarter = (warp - 2) >> 2;
        const int row = 32 * lg + lane;
        const int ty = row >> 3, tx = row & 7;
        QuantConsts q0 = {}, q1 = {};
        if (!kF32) {
            q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
            q1 = q0;
            if (a.y[1]) q1 = consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f);
        }
        const float fa = __fmul_rn(sx, a.s_w), fr = __fmul_rn(sx, a.c_w);
        const float floor_v = a.relu ? 0.0f : -INFINITY;
        int acc = 0; uint32_t aphase = 0;
        int hb = 0; uint32_t hphase = 0;
        for (int64_t tile = tile0; tile < tile_end; tile += tstep) {
            int n, oh0, ow0;
            tile_origin(tile, n, oh0, ow0);
            mbar_wait(hfull0 + 8 * hb, hphase);                   // the loaders' code sums of this halo are in S_s
            const int oh = oh0 + ty, ow = ow0 + tx;
            const bool valid = tile < a.tiles && oh < a.H && ow < a.W;      // (tile >= tiles: a ghost iteration)
            int suma = 0;
            uint32_t oob = 0;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                suma += S_s[hb][(ty + t / 3) * kHX + tx + t % 3];
                const int ih = oh + t / 3 - 1, iw = ow + t % 3 - 1;
                if (!((unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W)) oob |= 1u << t;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(hempty0 + 8 * hb);          // S_s[hb] (and, for this warp, the halo) may be reused
            const float row_c = __fmul_rn(fr, (float)suma);
            const int64_t g = ((int64_t)n * a.H + oh) * a.W + ow;
            mbar_wait(tfull0 + 8 * acc, aphase);
            tc_fence_after();
            const uint32_t tbase = tmem + ((uint32_t)(32 * lg) << 16) + (uint32_t)(acc * a.npass);
            float lo = INFINITY, hi = -INFINITY;
            for (int chunk = quarter; chunk * 16 < n_real; chunk += kHEpiWarps / 4) {
                uint32_t d[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
                             "%13, %14, %15}, [%16];"
                             : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]),
                               "=r"(d[8]), "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]),
                               "=r"(d[15])
                             : "r"(tbase + (uint32_t)(chunk * 16)));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (valid) {
                    const int nv = min(16, n_real - chunk * 16);
                    float yv[16];
                    if (oob) {                                     // border pixel: take the absent taps back out
                        float corr[16];
#pragma unroll
                        for (int j = 0; j < 16; ++j) corr[j] = 0.0f;
                        for (uint32_t mm = oob; mm; mm &= mm - 1) {
                            const float4 *tc = reinterpret_cast<const float4 *>(tap_c + (__ffs(mm) - 1) * a.npass + chunk * 16);
#pragma unroll
                            for (int j4 = 0; j4 < 4; ++j4) {
                                const float4 v4 = tc[j4];
                                corr[4 * j4] = __fadd_rn(corr[4 * j4], v4.x); corr[4 * j4 + 1] = __fadd_rn(corr[4 * j4 + 1], v4.y);
                                corr[4 * j4 + 2] = __fadd_rn(corr[4 * j4 + 2], v4.z); corr[4 * j4 + 3] = __fadd_rn(corr[4 * j4 + 3], v4.w);
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            yv[j] = fmaxf(__fmaf_rn(fa, (float)(int)d[j],
                                                    __fsub_rn(__fadd_rn(row_c, chan_c[chunk * 16 + j]), corr[j])), floor_v);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            yv[j] = fmaxf(__fmaf_rn(fa, (float)(int)d[j], __fadd_rn(row_c, chan_c[chunk * 16 + j])), floor_v);
                    }
                    if (nv < 16) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) yv[j] = j < nv ? yv[j] : 0.0f;
                    }
                    if (kF32) {
                        float *py = a.yf + (((int64_t)n * a.O + o0 + chunk * 16) * a.H + oh) * a.W + ow;
                        const int64_t plane = (int64_t)a.H * a.W;
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            if (j < nv) {
                                py[j * plane] = yv[j];
                                lo = fminf(lo, yv[j]); hi = fmaxf(hi, yv[j]);
                            }
                        }
                    } else {
                        const int64_t off = g * a.cp_out + o0 + chunk * 16;
                        for (int gi = 0; gi < 2; ++gi) {
                            if (!a.y[gi]) break;
                            uint32_t w[4];
                            pack_codes<16>(yv, gi ? q1 : q0, w);
                            if (nv < 16) {
#pragma unroll
                                for (int q4 = 0; q4 < 4; ++q4) w[q4] = keep_left(w[q4], nv - 4 * q4);
                            }
                            *reinterpret_cast<uint4 *>(a.y[gi] + off) = make_uint4(w[0], w[1], w[2], w[3]);
                        }
                    }
                }
            }
            if (kF32 && a.stats) {                                  // per-sample extrema, encoded (a tile is one image)
                lo = warp_min(lo); hi = warp_max(hi);
                if (lane == 0 && hi >= lo) {
                    atomicMax(a.stats + 2 * n, enc_ordered(-lo));
                    atomicMax(a.stats + 2 * n + 1, enc_ordered(hi));
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
            if (++acc == 2) { acc = 0; aphase ^= 1; }
            hb ^= 1; if (hb == 0) hphase ^= 1;
        }
    }
    else {
        // ===== halo loaders: thread p < 180 owns halo pixel p of every tile
        const int p = tid - 32 * (2 + kHEpiWarps);
        const int hy = p / kHX, hx = p - hy * kHX;
        const int Cp16 = (a.C + 15) / 16;                          // real 16-byte chunks of a pixel
        int hb = 0; uint32_t hphase = 0;
        for (int64_t tile = tile0; tile < tile_end; tile += tstep) {
            int n, oh0, ow0;
            tile_origin(tile, n, oh0, ow0);
            mbar_wait(hempty0 + 8 * hb, hphase ^ 1);
            if (p < kHY * kHX) {
                const int ih = oh0 - 1 + hy, iw = ow0 - 1 + hx;
                const bool inb = tile < a.tiles && (unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W;
                const uint4 *src = reinterpret_cast<const uint4 *>(a.x + (((int64_t)n * a.H + ih) * a.W + iw) * (Cp16 * 16));
                uint8_t *dst = smem + hb * halo_bytes + p * 16;
                int sacc = 0;
                for (int c0 = 0; c0 < a.cpt; c0 += 4) {            // four 128-bit loads in flight
                    uint4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        v[u] = (inb && c0 + u < Cp16) ? __ldg(src + c0 + u) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        *reinterpret_cast<uint4 *>(dst + (c0 + u) * kHChunk) = v[u];
                        sacc = __dp4a((int)v[u].x, 0x01010101, sacc); sacc = __dp4a((int)v[u].y, 0x01010101, sacc);
                        sacc = __dp4a((int)v[u].z, 0x01010101, sacc); sacc = __dp4a((int)v[u].w, 0x01010101, sacc);
                    }
                }
                S_s[hb][p] = sacc;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic stores -> the tensor core's proxy
            __syncwarp();
            if (lane == 0) mbar_arrive(hfull0 + 8 * hb);
            hb ^= 1; if (hb == 0) hphase ^= 1;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();                                              // the peer may still multicast into / arrive on this CTA
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)a.tmem_cols)
                     : "memory");
}

}  // namespace
}  // namespace dfq

using namespace dfq;

// chunks per tap of the packed reduction: the 16-byte channel chunks rounded up to a multiple of 4
extern "C" int dfq_conv_q8_halo_k_host(int C) { return 9 * (((C + 15) / 16 + 3) / 4 * 4) * 16; }

namespace dfq {
static int halo_launch(HaloArgs &a, int o_cap, bool f32, void *stream) {
    const int Cp = (a.C + 15) / 16 * 16;
    a.cpt = ((Cp / 16) + 3) / 4 * 4;
    const int kq = 9 * a.cpt * 16;
    int64_t bytes;
    int e = dfq_pw_i8_plan_cap_host(a.O, kq, o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    if (a.slabs != 9 * a.cpt / 4) return (int)cudaErrorInvalidValue;
    a.cp_out = (a.O + 15) / 16 * 16;
    a.tiles_x = (a.W + kHTX - 1) / kHTX; a.tiles_y = (a.H + kHTY - 1) / kHTY;
    a.tiles = (int64_t)a.N * a.tiles_x * a.tiles_y;
    if (a.tiles > 0x7fffffffll || (int64_t)a.N * a.H * a.W > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    a.tmem_cols = 32;
    while (a.tmem_cols < 2 * a.npass) a.tmem_cols *= 2;
    if (a.tmem_cols > 512) return (int)cudaErrorInvalidValue;
    const size_t halo_bytes = (size_t)a.cpt * kHChunk, bslab = (size_t)a.npass * kHSlab;
    const size_t tap_bytes = 9 * (size_t)a.npass * sizeof(float);
    const size_t smem_cap = 227 * 1024 - 4096;                 // the kernel has ~2.7 KB of static shared memory
    if (2 * halo_bytes + 4 * bslab + tap_bytes + 128 > smem_cap) return (int)cudaErrorInvalidValue;
    int stages = (int)((smem_cap - 2 * halo_bytes - tap_bytes - 128) / (2 * bslab));   // a stage = two slabs
    if (stages > kHMaxStages) stages = kHMaxStages;
    if (stages < 2 || a.slabs > 128) return (int)cudaErrorInvalidValue;
    a.stages = stages;
    const size_t smem = 2 * halo_bytes + (size_t)stages * 2 * bslab + tap_bytes + 128;
    static bool configured = false;
    if (!configured) {
        cudaError_t ce = cudaFuncSetAttribute(conv3x3_halo_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
        if (ce == cudaSuccess)
            ce = cudaFuncSetAttribute(conv3x3_halo_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
        if (ce != cudaSuccess) return (int)ce;
        configured = true;
    }
    // clusters of two CTAs, one CTA per SM (two accumulators = all TMEM columns at N = 256); pairs share a pass
    int64_t pairs_per_pass = (int64_t)(sm_count() / 2) / a.n_pass;
    if (pairs_per_pass < 1) pairs_per_pass = 1;
    if (pairs_per_pass > (a.tiles + 1) / 2) pairs_per_pass = (a.tiles + 1) / 2;
    const unsigned grid = (unsigned)(2 * pairs_per_pass * a.n_pass);
    if (f32) conv3x3_halo_kernel<true><<<grid, kHThreads, smem, (cudaStream_t)stream>>>(a);
    else conv3x3_halo_kernel<false><<<grid, kHThreads, smem, (cudaStream_t)stream>>>(a);
    g_launches += 1;
    return (int)cudaGetLastError();
}

// the shape both forms cover; `smem_ok`: two halo buffers + at least two weight slabs fit
bool conv_halo_covers(int C, int O, int kh, int kw, int stride, int pad, int dil) {
    if (kh != 3 || kw != 3 || stride != 1 || dil != 1 || pad != 1) return false;
    const int cpt = (((C + 15) / 16) + 3) / 4 * 4;
    const int npass = O >= 256 ? 256 : (O + 15) / 16 * 16;
    return 2 * (size_t)cpt * kHChunk + 4 * (size_t)npass * kHSlab + 9 * (size_t)npass * 4 + 128 <= 227 * 1024 - 4096;
}

int conv_q8_halo(const DfqConvQ8Desc *d, void *stream) {
    if (!conv_halo_covers(d->C, d->O, d->kh, d->kw, d->stride, d->pad, d->dil) || d->OH != d->H || d->OW != d->W)
        return (int)cudaErrorInvalidValue;
    HaloArgs a = {};
    a.x = d->x; a.wpk = d->wpk; a.wr = d->wr; a.wrt = d->wrt; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.out_slot[0] = (const RangeSlot *)d->out_slot0; a.out_slot[1] = (const RangeSlot *)d->out_slot1;
    a.y[0] = d->y0; a.y[1] = d->y1;
    a.N = d->N; a.C = d->C; a.H = d->H; a.W = d->W; a.O = d->O; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    return halo_launch(a, d->o_cap, false, stream);
}

// dynamic-range plan (dfq_conv_i8 with halo != 0): the input already quantised into `codes` by the pre-pass
int conv_i8_halo(const DfqConvI8Desc *d, const int8_t *codes, void *stream) {
    if (!conv_halo_covers(d->C, d->O, d->kh, d->kw, d->stride, d->pad, d->dil) || d->OH != d->H || d->OW != d->W ||
        d->bits != 8 || !codes)
        return (int)cudaErrorInvalidValue;
    HaloArgs a = {};
    a.x = codes; a.wpk = d->wpk; a.wr = d->wr; a.wrt = d->wrt; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.yf = d->y; a.stats = d->out_stats;
    a.N = d->N; a.C = d->C; a.H = d->H; a.W = d->W; a.O = d->O; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    return halo_launch(a, d->o_cap, true, stream);
}
}  // namespace dfq
