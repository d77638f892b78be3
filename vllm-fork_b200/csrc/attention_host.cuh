// b200ops -- paged-attention decode: host-side planning, TMA descriptors, launch (included by attention.cu).
#pragma once

#include <mutex>
#include <unordered_map>

namespace b200 {
namespace attn {

constexpr int kMaxCtas = 192;  // workspace is sized for at most this many persistent CTAs

struct TmapKey {
  const void* base;
  int64_t nb, sb, st, sh;
  int bs, hkv, hc, D, es, kv_pair;
  bool operator==(const TmapKey& o) const {
    return base == o.base && nb == o.nb && sb == o.sb && st == o.st && sh == o.sh && bs == o.bs &&
           hkv == o.hkv && hc == o.hc && D == o.D && es == o.es && kv_pair == o.kv_pair;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = (size_t)k.base * 1315423911u;
    h ^= (size_t)k.nb * 2654435761u + (size_t)k.sb * 40503u + (size_t)k.st * 7919u +
         (size_t)k.sh * 104729u + (size_t)(k.bs * 31 + k.hkv * 17 + k.hc * 13 + k.D * 7 + k.es + k.kv_pair * 3);
    return h;
  }
};
static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmaps;
static std::mutex g_tmap_mu;

// 5-D view of a paged K (or V) cache: [inner 128 B][token in block][128-B segment][head][block].
// kv_pair (K|V packed per token and head, one 128-byte segment per row: the fork's NHD layout with FP8 D = 128):
// the idle segment dimension becomes the K / V selector (stride = one row), so ONE box brings the K tile and the V
// tile of a 16-token page: [128 B][token][K, V][head][block] lands as [K tile | V tile] - the same shared-memory
// image as the two separate loads when the CTA handles one head (hc == 1).
static int get_tmap(CUtensorMap* out, const void* base, int64_t nb, int64_t sb, int64_t st,
                    int64_t sh, int bs, int hkv, int hc, int D, int es, int kv_pair = 0) {
  TmapKey key{base, nb, sb, st, sh, bs, hkv, hc, D, es, kv_pair};
  std::lock_guard<std::mutex> lk(g_tmap_mu);
  auto it = g_tmaps.find(key);
  if (it != g_tmaps.end()) {
    *out = it->second;
    return B200_OK;
  }
  const int RB = D * es;
  const int IB = RB < 128 ? RB : 128;
  const int nseg = RB / IB;
  uint64_t dims[5] = {(uint64_t)(IB / es), (uint64_t)bs, (uint64_t)nseg, (uint64_t)hkv, (uint64_t)nb};
  uint64_t strides[4] = {(uint64_t)(st * es), (uint64_t)IB, (uint64_t)(sh * es), (uint64_t)(sb * es)};
  uint32_t box[5] = {(uint32_t)(IB / es), 16u, (uint32_t)nseg, (uint32_t)hc, 1u};
  if (kv_pair) {   // nseg == 1: dimension 2 selects K (0) or V (1), one row = RB bytes apart
    dims[2] = 2;
    strides[1] = (uint64_t)RB;
    box[2] = 2;
  }
  CUtensorMap tm;
  int rc = encode_tensor_map(&tm, es == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16,
                             5, base, dims, strides, box,
                             IB == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  if (rc) return rc;
  if (g_tmaps.size() > 8192) g_tmaps.clear();
  g_tmaps.emplace(key, tm);
  *out = tm;
  return B200_OK;
}

struct Plan {
  int hc, tg, n_chunks, stages, nt, grid_x;
  uint32_t tile_bytes, stage_bytes, off_merge, off_cum, off_bar, smem_bytes;
};

static int make_plan(Plan& pl, int S, int Hq, int Hkv, int D, int es, int sms, int smem_optin) {
  const int g = Hq / Hkv;
  const int RB = D * es;
  pl.nt = (g + 7) / 8;
  int hc = 1;
  for (int c = 1; c <= NW && c <= Hkv; c++)
    if (Hkv % c == 0 && 2 * c * 16 * RB <= 65536) hc = c;
  int tg = NW / hc;
  auto merge_bytes = [&](int t) { return (t - 1) * hc * pl.nt * (64 + 32 * D); };
  while (tg > 1 && (tg * 2 * hc * 16 * RB > 65536 || merge_bytes(tg) > 32768)) tg--;
  pl.hc = hc;
  pl.tg = tg;
  pl.n_chunks = Hkv / hc;
  pl.tile_bytes = (uint32_t)(hc * 16 * RB);
  pl.stage_bytes = (uint32_t)tg * 2u * pl.tile_bytes;
  const uint32_t mb = (uint32_t)merge_bytes(tg);
  const uint32_t cb = (uint32_t)(((S + 1) * 4 + 15) & ~15);
  const uint32_t bb = 2u * 8u * MAX_STAGES;
  const int64_t avail = (int64_t)smem_optin - 1024 - 64 /*static*/ - mb - cb - bb;
  int stages = (int)(avail / pl.stage_bytes);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) return B200_ERR_UNSUPPORTED;
  pl.stages = stages;
  pl.off_merge = (uint32_t)stages * pl.stage_bytes;
  pl.off_cum = pl.off_merge + mb;
  pl.off_bar = pl.off_cum + cb;
  pl.smem_bytes = pl.off_bar + bb + 1024;
  pl.grid_x = sms / pl.n_chunks;
  if (pl.grid_x < 1) pl.grid_x = 1;
  if (pl.grid_x * pl.n_chunks > kMaxCtas) pl.grid_x = kMaxCtas / pl.n_chunks;
  if (pl.grid_x < 1) return B200_ERR_UNSUPPORTED;
  return B200_OK;
}

static size_t ws_cum_bytes(int S) { return (size_t)(((S + 2) * 4 + 255) & ~255); }

template <int D, int NT, int MODE>
static int launch(const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, const Plan& pl,
                  cudaStream_t stream) {
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  static int configured_smem = 0;
  auto kern = paged_attn_decode_kernel<D, NT, MODE>;
  if (!configured || configured_smem < (int)pl.smem_bytes) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)pl.smem_bytes));
    configured = true;
    configured_smem = (int)pl.smem_bytes;
  }
  dim3 grid(pl.grid_x, pl.n_chunks);
  B200_CUDA(launch_pdl(kern, grid, dim3(NTHREADS), (size_t)pl.smem_bytes, stream, tmK, tmV, p));
  dim3 rgrid(p.S, pl.n_chunks);
  constexpr bool QBF16 = (MODE == BF16_BF16 || MODE == BF16_FP8);
  int rthreads = p.hc * p.g * (D / 4);
  rthreads = rthreads > 1024 ? 1024 : ((rthreads + 31) / 32) * 32;
  // launched with the programmatic attribute: its CTAs are resident and waiting when the decode kernel drains
  B200_CUDA(launch_pdl(paged_attn_reduce_kernel<QBF16>, rgrid, dim3(rthreads), 0, stream, p, D, pl.grid_x));
  return B200_OK;
}

}  // namespace attn
}  // namespace b200

extern "C" size_t b200_paged_attention_workspace_bytes(int num_seqs, int num_heads,
                                                       int num_kv_heads, int head_size) {
  (void)num_kv_heads;
  using namespace b200::attn;
  if (num_seqs < 0 || num_heads <= 0 || head_size <= 0) return 0;
  return ws_cum_bytes(num_seqs) + (size_t)kMaxCtas * 2 * num_heads * (head_size + 2) * sizeof(float);
}

extern "C" int b200_paged_attention_decode(const b200_paged_attn_args* a, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::attn;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(a != nullptr, B200_ERR_INVALID, "paged_attention_decode: null args");
  const int S = a->num_seqs, Hq = a->num_heads, Hkv = a->num_kv_heads, D = a->head_size;
  B200_CHECK(S >= 0 && Hq > 0 && Hkv > 0 && D > 0, B200_ERR_INVALID,
             "paged_attention_decode: bad sizes (seqs %d heads %d kv_heads %d head_size %d)", S, Hq,
             Hkv, D);
  if (S == 0) return B200_OK;
  B200_CHECK(Hq % Hkv == 0, B200_ERR_INVALID,
             "paged_attention_decode: num_heads (%d) must be divisible by num_kv_heads (%d)", Hq, Hkv);
  B200_CHECK(a->out && a->q && a->k_cache && a->v_cache && a->block_table && a->seq_lens &&
                 a->workspace,
             B200_ERR_INVALID, "paged_attention_decode: null pointer");
  B200_CHECK(a->block_size > 0 && a->block_size % 16 == 0, B200_ERR_INVALID,
             "paged_attention_decode: block_size (%d) must be a multiple of 16", a->block_size);
  B200_CHECK(a->q_dtype == B200_BF16 || a->q_dtype == B200_F16, B200_ERR_INVALID,
             "paged_attention_decode: query dtype must be bf16 or f16");
  const bool fp8 = a->kv_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8 || a->kv_dtype == a->q_dtype, B200_ERR_INVALID,
             "paged_attention_decode: kv dtype must equal query dtype (\"auto\") or be fp8_e4m3");
  const int es = fp8 ? 1 : 2;
  B200_CHECK(D == 64 || D == 128 || D == 256, B200_ERR_UNSUPPORTED,
             "paged_attention_decode: head_size %d not supported (64, 128, 256)", D);
  const int g = Hq / Hkv;
  B200_CHECK(g <= 16, B200_ERR_UNSUPPORTED,
             "paged_attention_decode: more than 16 query heads per kv head (%d) not supported", g);
  B200_CHECK(S <= 16384, B200_ERR_UNSUPPORTED,
             "paged_attention_decode: more than 16384 sequences per call (%d)", S);
  auto al16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0; };
  B200_CHECK(al16(a->q) && al16(a->out) && al16(a->k_cache) && al16(a->v_cache) &&
                 (a->q_stride_s * 2) % 16 == 0 && (a->q_stride_h * 2) % 16 == 0 &&
                 (a->out_stride_s * 2) % 16 == 0 && (a->out_stride_h * 2) % 16 == 0,
             B200_ERR_INVALID, "paged_attention_decode: q/out/cache pointers and strides must be 16-byte aligned");
  B200_CHECK((a->kv_block_stride * es) % 16 == 0 && (a->kv_token_stride * es) % 16 == 0 &&
                 (a->kv_head_stride * es) % 16 == 0 && a->kv_block_stride > 0 &&
                 a->kv_token_stride > 0 && a->kv_head_stride > 0,
             B200_ERR_INVALID, "paged_attention_decode: kv cache strides must be positive multiples of 16 bytes");
  B200_CHECK(a->num_blocks > 0, B200_ERR_INVALID, "paged_attention_decode: num_blocks must be > 0");
  const int sms = sm_count();
  const int smem = max_smem_optin();
  B200_CHECK(sms > 0 && smem > 0, B200_ERR_CUDA,
             "paged_attention_decode: no CUDA device (b200ops has no CPU fallback)");
  Plan pl;
  int rc = make_plan(pl, S, Hq, Hkv, D, es, sms, smem);
  B200_CHECK(rc == B200_OK, rc, "paged_attention_decode: no shared-memory plan for this shape");
  const size_t need = b200_paged_attention_workspace_bytes(S, Hq, Hkv, D);
  B200_CHECK(a->workspace_bytes >= need, B200_ERR_WORKSPACE,
             "paged_attention_decode: workspace too small (%zu < %zu bytes)", a->workspace_bytes, need);
  B200_CHECK((reinterpret_cast<uintptr_t>(a->workspace) & 255) == 0, B200_ERR_INVALID,
             "paged_attention_decode: workspace must be 256-byte aligned");

  CUtensorMap tmK, tmV;
  // K|V packed per (token, head) with one 128-byte row each and a single head per CTA (TP shards with one kv head):
  // one TMA box per page instead of two (the producer is bound by TMA issue there: 2 KB per instruction)
  const int kv_pair = (pl.hc == 1 && D * es <= 128 &&
                       reinterpret_cast<const uint8_t*>(a->v_cache) == reinterpret_cast<const uint8_t*>(a->k_cache) + D * es &&
                       a->kv_head_stride >= 2 * D)
                          ? 1 : 0;
  rc = get_tmap(&tmK, a->k_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride,
                a->kv_head_stride, a->block_size, Hkv, pl.hc, D, es, kv_pair);
  if (rc) return rc;
  rc = get_tmap(&tmV, a->v_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride,
                a->kv_head_stride, a->block_size, Hkv, pl.hc, D, es);
  if (rc) return rc;

  Params p;
  p.out = a->out;
  p.lse = a->lse;
  p.q = a->q;
  p.block_table = a->block_table;
  p.seq_lens = a->seq_lens;
  p.k_scale = fp8 ? a->k_scale : nullptr;
  p.v_scale = fp8 ? a->v_scale : nullptr;
  p.alibi = a->alibi_slopes;
  p.out_scale = a->output_scale;
  uint8_t* ws = reinterpret_cast<uint8_t*>(a->workspace);
  p.cum = reinterpret_cast<int32_t*>(ws);
  const int n_cta = pl.grid_x * pl.n_chunks;
  const int hq_chunk = pl.hc * g;
  p.part_ml = reinterpret_cast<float*>(ws + ws_cum_bytes(S));
  p.part_acc = p.part_ml + (size_t)n_cta * 2 * hq_chunk * 2;
  p.q_stride_s = a->q_stride_s;
  p.q_stride_h = a->q_stride_h;
  p.out_stride_s = a->out_stride_s;
  p.out_stride_h = a->out_stride_h;
  p.bt_stride = a->bt_stride;
  p.S = S; p.Hq = Hq; p.Hkv = Hkv; p.g = g; p.bs = a->block_size;
  p.hc = pl.hc; p.tg = pl.tg; p.stages = pl.stages; p.min_tpc = pl.tg;
  p.window = a->sliding_window > 0 ? a->sliding_window : 0;
  p.scale = a->scale;
  p.softcap = a->softcap > 0.f ? a->softcap : 0.f;
  p.tile_bytes = pl.tile_bytes;
  p.kv_pair = kv_pair;
  p.early_plan = pdl_level() >= 2 ? 1 : 0;
  p.off_merge = pl.off_merge; p.off_cum = pl.off_cum; p.off_bar = pl.off_bar;

  const int mode = fp8 ? (a->q_dtype == B200_BF16 ? BF16_FP8 : F16_FP8)
                       : (a->q_dtype == B200_BF16 ? BF16_BF16 : F16_F16);
#define B200_ATTN_CASE(DD, NN, MM) \
  if (D == DD && pl.nt == NN && mode == MM) return launch<DD, NN, MM>(tmK, tmV, p, pl, stream);
#define B200_ATTN_D(DD)               \
  B200_ATTN_CASE(DD, 1, BF16_BF16)    \
  B200_ATTN_CASE(DD, 2, BF16_BF16)    \
  B200_ATTN_CASE(DD, 1, F16_F16)      \
  B200_ATTN_CASE(DD, 2, F16_F16)      \
  B200_ATTN_CASE(DD, 1, BF16_FP8)     \
  B200_ATTN_CASE(DD, 2, BF16_FP8)     \
  B200_ATTN_CASE(DD, 1, F16_FP8)      \
  B200_ATTN_CASE(DD, 2, F16_FP8)
  B200_ATTN_D(64)
  B200_ATTN_D(128)
  B200_ATTN_D(256)
#undef B200_ATTN_D
#undef B200_ATTN_CASE
  set_last_error("paged_attention_decode: no kernel for head_size %d, group %d, mode %d", D, g, mode);
  return B200_ERR_UNSUPPORTED;
}
