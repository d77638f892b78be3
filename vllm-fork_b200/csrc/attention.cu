// b200ops -- paged-attention decode for sm_100a (B200).
//
// Contract: decode branch of AttentionImpl.forward (vllm/v1/attention/backend.py:895-934) as
// implemented by the reference's Triton unified attention (ops/triton_unified_attention.py:179-772):
// one query token per sequence, paged KV addressed through block_table, GQA, bf16/f16 or FP8-E4M3
// KV with per-tensor k/v scales, optional ALiBi / sliding window / soft-cap, fp32 softmax.
//
// B200 design:
//  * HBM-bound stream of K/V pages. One persistent CTA per SM; the flattened list of 16-token KV
//    tiles of ALL sequences is cut into equal contiguous ranges, one per CTA (perfect balance for
//    ragged batches, split-KV for long contexts falls out for free). A sequence cut by a CTA
//    boundary leaves an fp32 (m, l, acc) partial in the workspace; a small reduce kernel merges.
//  * A producer warp streams tiles with TMA (cp.async.bulk.tensor.5d, SWIZZLE_128B) into a deep
//    mbarrier ring (up to ~200 KB in flight per SM). The 5-D tensor map is built from the cache's
//    strides, so NHD / HND / split-K-V layouts are all the same code. Inside smem the tile is
//    [head][segment][token][128 B] so that every later shared-memory read is conflict-free.
//  * 8 consumer warps; warp <-> (kv head, token group). Per 16-token tile a warp does
//    S^T[16 tok x 8 q-heads] = K.Q^T and O^T[D x 8 q-heads] += V^T.P^T with mma.sync m16n8k16
//    (f16 for FP8 KV -- e4m3 -> f16 is exact -- bf16 otherwise), fp32 accumulation, tokens on the
//    M axis so a GQA group of <= 8 q-heads shares every K/V byte read from smem exactly once.
//    At FP8/GQA-4 the op needs ~52 TFLOP/s of useful math at HBM speed -- above what the fp32 FMA
//    pipe sustains under the power cap -- hence tensor cores for the two small GEMMs.
//    The contraction index of both GEMMs is permuted so each lane reads 16-byte chunks straight
//    from the swizzled TMA image (no ldmatrix, no transposes through smem); V's token pairing is
//    done with PRMT; P^T is produced from S^T with movmatrix.
//  * Online softmax in the exp2 domain with warp-shuffle max / sum reductions; the running max is
//    only re-based when it grows by more than 2^8 (same idea as the reference CPU kernel,
//    csrc/cpu/cpu_attn_impl.hpp:1288-1297), which keeps the rescale off the common path.
#include "common.cuh"

namespace b200 {
namespace attn {

constexpr int NW = 8;                       // consumer warps
constexpr int NTHREADS = 32 * (NW + 1);     // + 1 producer warp
constexpr int MAX_STAGES = 12;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float RESCALE_THRESHOLD = 8.0f;   // log2 units

enum Mode { BF16_BF16 = 0, F16_F16 = 1, BF16_FP8 = 2, F16_FP8 = 3 };

struct Params {
  void* out;
  float* lse;
  const void* q;
  const int32_t* block_table;
  const int32_t* seq_lens;
  const float* k_scale;
  const float* v_scale;
  const float* out_scale;   // nullable: fused static FP8 output
  const float* alibi;
  int32_t* cum;      // workspace: [S + 2] tile prefix sums (+ tiles-per-CTA at [S+1])
  float* part_acc;   // workspace: [n_cta * 2][hq_chunk][D]
  float* part_ml;    // workspace: [n_cta * 2][hq_chunk][2]
  int64_t q_stride_s, q_stride_h, out_stride_s, out_stride_h, bt_stride;
  int S, Hq, Hkv, g, bs;
  int hc, tg, stages, min_tpc;
  int window;
  float scale, softcap;
  uint32_t tile_bytes;   // hc * 16 * RB  (one K or V tile)
  int early_plan;        // PDL level 2: the tile plan (reads seq_lens only) precedes griddepcontrol.wait
  int kv_pair;           // 1: tmK's box carries the K tile and the V tile of a page (attention_host.cuh get_tmap)
  uint32_t off_merge, off_cum, off_bar;  // smem offsets
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(a));
  return r;
}
template <bool BF16>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4],
                                         const uint32_t (&b)[2]) {
  if constexpr (BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
}
template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (BF16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
__device__ __forceinline__ uint32_t bf16x2_to_f16x2(uint32_t v) {
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&v);
  float2 f = __bfloat1622float2(b);
  __half2 h = __floats2half2_rn(f.x, f.y);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(addr));
  return r;
}
__device__ __forceinline__ void bar_sync_named(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Compile-time geometry of one (head_size, kv element size) combination.
template <int D, int ES>
struct Geo {
  static constexpr int RB = D * ES;                    // bytes of one K (or V) row
  static constexpr int IB = RB < 128 ? RB : 128;       // TMA inner box bytes == swizzle span
  static constexpr int NSEG = RB / IB;                 // 128-byte segments per row
  static constexpr int EPC = 16 / ES;                  // elements per 16-byte chunk
  static constexpr int NCK = RB / 64;                  // K chunks per lane per row
  static constexpr int KS = D / 16;                    // QK k-steps
  static constexpr int MT = D / 16;                    // PV m-tiles
  static constexpr int VB = RB / 8;                    // V bytes per (lane row, token)
  static constexpr int NCV = VB >= 16 ? VB / 16 : 1;   // V chunks per (lane row, token)
  static constexpr int SWZ = IB == 128 ? 7 : 3;        // swizzle mask (128B / 64B)
  static_assert(RB == 64 || RB == 128 || RB == 256 || RB == 512, "unsupported head_size");
};

// byte offset of (row-in-tile-slice, 16B chunk) inside one [16 tok][IB] slice, TMA swizzle applied
template <int IB>
__device__ __forceinline__ uint32_t swz_off(int tok, int chunk) {
  if constexpr (IB == 128)
    return (uint32_t)(tok * 128 + ((chunk ^ (tok & 7)) << 4));
  else
    return (uint32_t)(tok * 64 + ((chunk ^ ((tok >> 1) & 3)) << 4));
}

}  // namespace attn
}  // namespace b200

#include "attention_kernel.cuh"
#include "attention_host.cuh"

// ------------------------------------------------------------------------------------------------
// b200_paged_attention_decode_static: paged-attention decode for callers that cannot allocate.
//
// Who that is: code that runs where no CUDA allocator may be invoked. Two kinds of caller are meant.
// A serving worker whose device memory was measured and handed out at start-up -- whatever the
// weights left over became KV-cache blocks -- has no headroom afterwards: a later cudaMalloc fails,
// and it also synchronises the whole device in the middle of a decode step. And a binding written
// against this C ABI from a runtime that owns no device allocator (a bare cgo / JNI / ctypes caller
// that only holds the buffers it was handed) cannot produce a workspace pointer at all. The regular
// entry point asks both for a split-KV workspace they have no means to provide; here the workspace
// is a fixed array that is part of the library image.
//
// What it costs: B200_PAGED_ATTN_STATIC_WS_BYTES (8 MiB) of device memory on every GPU that loads
// libb200ops.so, held until the process exits, whether or not this entry point is ever called.
//
// g_static_ws is only a device address inside device code. On the host the same name is the
// runtime's registration handle for the symbol, so the address the kernels need is looked up
// with cudaGetSymbolAddress -- once per device, with that device current, in static_ws_address().
namespace b200 {
namespace attn {

constexpr size_t kStaticWsBytes = B200_PAGED_ATTN_STATIC_WS_BYTES;
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

static void* g_static_ws_addr[64] = {};   // device address of g_static_ws, per device ordinal
static std::mutex g_static_ws_mu;

// Device address of the scratch array on the CURRENT device. The first call on a device resolves
// it (this may also make the runtime load the library's device code there); later calls are a
// table read. Never allocates.
static int static_ws_address(void** out) {
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_CHECK(dev >= 0 && dev < 64, B200_ERR_UNSUPPORTED,
             "paged_attention_decode_static: device ordinal %d outside the supported range 0..63", dev);
  std::lock_guard<std::mutex> lk(g_static_ws_mu);
  if (g_static_ws_addr[dev] == nullptr) {
    void* ptr = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&ptr, g_static_ws));
    B200_CHECK(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0, B200_ERR_CUDA,
               "paged_attention_decode_static: the runtime returned an unusable scratch address (%p)", ptr);
    g_static_ws_addr[dev] = ptr;
  }
  *out = g_static_ws_addr[dev];
  return B200_OK;
}

}  // namespace attn
}  // namespace b200

extern "C" int b200_paged_attention_decode_static(const b200_paged_attn_args* a, b200_stream_t stream) {
  using namespace b200;
  using namespace b200::attn;
  B200_CHECK(a != nullptr, B200_ERR_INVALID, "paged_attention_decode_static: null args");
  // Size gate first: it needs no device, and an oversized request must never reach the kernels.
  // (Nonsensical sizes make the query return 0; the regular entry point below names what is wrong.)
  const size_t need = b200_paged_attention_workspace_bytes(a->num_seqs, a->num_heads, a->num_kv_heads,
                                                           a->head_size);
  B200_CHECK(need <= kStaticWsBytes, B200_ERR_WORKSPACE,
             "paged_attention_decode_static: %d sequences x %d heads x head_size %d need %zu bytes of split-KV "
             "scratch, the built-in array holds %zu; use b200_paged_attention_decode with a caller-owned workspace",
             a->num_seqs, a->num_heads, a->head_size, need, kStaticWsBytes);
  if (a->num_seqs == 0) return b200_paged_attention_decode(a, stream);   // nothing to launch
  void* ws = nullptr;
  int rc = static_ws_address(&ws);
  if (rc != B200_OK) return rc;
  // Same validation, plan, kernels and launch order as the regular entry point; only the workspace
  // differs. Nothing in it has to start at zero: the decode kernel writes the tile prefix sums and
  // every partial it later asks the reduce kernel to read, so stale bytes from the previous call
  // (the array is shared by all callers and never cleared) are never consumed.
  b200_paged_attn_args args = *a;
  args.workspace = ws;
  args.workspace_bytes = kStaticWsBytes;
  return b200_paged_attention_decode(&args, stream);
}
