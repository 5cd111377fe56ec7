// Fused attention kernels: CrossAttention.forward's softmax(q k^T / sqrt(d)) v
// (lvdm/modules/attention.py:92-114) with head dim 64, without materialising the score matrix.
#pragma once
#include "common.cuh"

namespace moca {

// Row addressing of one "outer" index o (a frame for spatial attention, a (b, h*w) column for temporal attention):
//   first row = (o / div) * mul_a + (o % div) * mul_b ; consecutive sequence elements are `step` rows apart.
struct SeqMap {
    int div = 1, mul_a = 0, mul_b = 0, step = 1;
};

struct AttnArgs {
    const op_t* q = nullptr; int ldq = 0; SeqMap qmap;     // q row r of outer o, head h: q[(row)*ldq + h*64 ..]
    const op_t* k = nullptr; const op_t* v = nullptr; int ldkv = 0; SeqMap kvmap;
    op_t* out = nullptr; int ldo = 0;                       // addressed with qmap
    int n_outer = 0, heads = 0, Lq = 0, Lk = 0;
    int64_t q_rows = 0, kv_rows = 0;   // rows of the q / kv buffers (TMA bounds of the tcgen05 kernel; 0: unknown -> mma.sync kernel)
    const int* kv_len = nullptr;   // optional device int[n_outer / kv_len_div]: valid keys (cross-attention)
    int kv_len_div = 1;
    float scale = 0.125f;
};

// Flash-style kernel: 64 query rows per CTA, 64-key tiles, online softmax (spatial self- and cross-attention).
int attention_flash(const AttnArgs& a, cudaStream_t st);
// tcgen05 / TMEM / TMA flash kernel (attn_tc.cu): 128 query rows per CTA, 128-key tiles, S and P.V accumulated in TMEM.
// Needs contiguous sequences (step == 1), q_rows / kv_rows, Lq >= 64.  Planned once (tensor maps), run many times.
struct AttnTcLaunch {
    alignas(64) unsigned char storage[640];
};
bool attention_tc_supported(const AttnArgs& a);
int attention_tc_plan(const AttnArgs& a, AttnTcLaunch* out);
int attention_tc_run(const AttnTcLaunch& l, cudaStream_t st);
// One warp per sequence, Lq = Lk <= 16 (temporal attention over frames).
int attention_short(const AttnArgs& a, cudaStream_t st);

}  // namespace moca
