// The two steps either side of the model that SURVEY 8(f) ranks "next":
//   f2  input pipeline on the device: per-utterance zero-mean / unit-variance normalisation (what
//       Wav2Vec2Processor does at centralized/finetune_ASRs.py:88-96), zero padding to the longest utterance, attention
//       mask and conv-stack frame lengths (the collator at finetune_ASRs.py:59-86) from ONE packed host->device copy
//   f4  word error rate on the device: greedy ids -> words (split at the delimiter id) -> Levenshtein distance against the
//       reference transcript ids (federated/src/update.py:36-50 compute_metrics -> jiwer.wer)
// Both HBM / latency bound, deterministic (fixed-order reductions, integer DP).
#include "common.cuh"

namespace b200asr {

// ------------------------------------------------------------------------------------------------ f2 normalise + pad
constexpr int NORM_CHUNK = 8192;  // samples per block in the statistics pass

// partial[b][chunk] = (sum, sum of squares) of the chunk in double (the combine stays exact enough for 30 s utterances)
__global__ void __launch_bounds__(256) wave_stats_kernel(const float* __restrict__ packed, const long long* __restrict__ offsets,
                                                         double* __restrict__ partial, int max_chunks) {
    __shared__ double red[2][8];
    const int b = blockIdx.y, chunk = blockIdx.x;
    const long long o0 = offsets[b], len = offsets[b + 1] - o0;
    const long long c0 = (long long)chunk * NORM_CHUNK;
    double s = 0.0, q = 0.0;
    if (c0 < len) {
        const long long c1 = min(len, c0 + NORM_CHUNK);
        for (long long i = c0 + threadIdx.x; i < c1; i += blockDim.x) {
            const double v = (double)packed[o0 + i];
            s += v;
            q += v * v;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { red[0][w] = s; red[1][w] = q; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ts = 0.0, tq = 0.0;
        for (int i = 0; i < 8; ++i) { ts += red[0][i]; tq += red[1][i]; }
        partial[((long long)b * max_chunks + chunk) * 2] = ts;
        partial[((long long)b * max_chunks + chunk) * 2 + 1] = tq;
    }
}

// out[b][i] = (x - mean_b) / sqrt(var_b + 1e-7) for i < len_b, else 0; mask[b][i] = i < len_b; frames[b] = conv lengths
__global__ void __launch_bounds__(256) wave_normalize_pad_kernel(const float* __restrict__ packed,
                                                                 const long long* __restrict__ offsets,
                                                                 const double* __restrict__ partial, float* __restrict__ out,
                                                                 int* __restrict__ mask, int* __restrict__ frames, int Lmax,
                                                                 int max_chunks, const int* __restrict__ conv_k,
                                                                 const int* __restrict__ conv_s, int n_conv) {
    __shared__ float sh_mean, sh_rstd;
    const int b = blockIdx.y;
    const long long o0 = offsets[b], len = offsets[b + 1] - o0;
    if (threadIdx.x == 0) {
        double s = 0.0, q = 0.0;
        const int nch = (int)((len + NORM_CHUNK - 1) / NORM_CHUNK);
        for (int c = 0; c < nch; ++c) {  // fixed order: deterministic
            s += partial[((long long)b * max_chunks + c) * 2];
            q += partial[((long long)b * max_chunks + c) * 2 + 1];
        }
        const double n = len > 0 ? (double)len : 1.0;
        const double mean = s / n;
        double var = q / n - mean * mean;  // population variance (np.var), as Wav2Vec2FeatureExtractor
        if (var < 0.0) var = 0.0;
        sh_mean = (float)mean;
        sh_rstd = (float)(1.0 / sqrt(var + 1e-7));
        if (blockIdx.x == 0 && frames) {
            long long l = len;
            for (int i = 0; i < n_conv; ++i) l = (l >= conv_k[i]) ? (l - conv_k[i]) / conv_s[i] + 1 : 0;
            frames[b] = (int)l;
        }
    }
    __syncthreads();
    const float mean = sh_mean, rstd = sh_rstd;
    const long long i0 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const long long stride = (long long)gridDim.x * blockDim.x * 4;
    float* ob = out + (long long)b * Lmax;
    int* mb = mask ? mask + (long long)b * Lmax : nullptr;
    for (long long i = i0; i < Lmax; i += stride) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long e = i + j;
            if (e >= Lmax) break;
            const bool valid = e < len;
            ob[e] = valid ? (packed[o0 + e] - mean) * rstd : 0.f;
            if (mb) mb[e] = valid ? 1 : 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------ f4 words + edit distance
// ids [B, ld] (collapsed greedy ids or label ids; entries < 0 are padding and skipped), len[b] valid prefix (NULL = ld).
// A word = maximal run of ids != delim; its token is the FNV-1a hash of the ids.  delim < 0: every id is its own token
// (character error rate).  One thread per sequence: a few hundred sequential steps, negligible next to the model.
__global__ void words_from_ids_kernel(const int* __restrict__ ids, const int* __restrict__ len, int B, int ld, int delim,
                                      unsigned long long* __restrict__ words, int* __restrict__ nwords, int ldw) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int n = len ? min(max(len[b], 0), ld) : ld;
    const int* p = ids + (long long)b * ld;
    unsigned long long* w = words + (long long)b * ldw;
    int nw = 0;
    unsigned long long h = 1469598103934665603ULL;
    bool open = false;
    for (int i = 0; i < n; ++i) {
        const int v = p[i];
        if (v < 0) continue;
        if (delim < 0) {
            if (nw < ldw) w[nw] = (unsigned long long)(unsigned)v;
            ++nw;
        } else if (v == delim) {
            if (open) {
                if (nw < ldw) w[nw] = h;
                ++nw;
                open = false;
                h = 1469598103934665603ULL;
            }
        } else {
            h = (h ^ (unsigned long long)(unsigned)v) * 1099511628211ULL;
            open = true;
        }
    }
    if (open) {
        if (nw < ldw) w[nw] = h;
        ++nw;
    }
    nwords[b] = min(nw, ldw);
}

// Levenshtein distance between token rows a[b][:na] and r[b][:nr], one warp per pair, anti-diagonal wavefront: the cells
// of diagonal k = i + j depend on diagonals k-1 and k-2 only, so a warp sweeps a diagonal in parallel.  Three rotating
// diagonal buffers of (lda + 1) ints per warp in dynamic shared memory.
__global__ void __launch_bounds__(128) edit_distance_kernel(const unsigned long long* __restrict__ a, const int* __restrict__ a_len,
                                                            const unsigned long long* __restrict__ r, const int* __restrict__ r_len,
                                                            int* __restrict__ dist, int B, int lda, int ldr) {
    extern __shared__ int dsm[];
    const int wpb = blockDim.x >> 5, wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * wpb + wib;
    if (b >= B) return;
    int* d0 = dsm + wib * 3 * (lda + 1);  // diagonal k-2, indexed by i
    int* d1 = d0 + (lda + 1);             // diagonal k-1
    int* d2 = d1 + (lda + 1);             // diagonal k
    const int n = min(max(a_len[b], 0), lda), m = min(max(r_len[b], 0), ldr);
    const unsigned long long* ap = a + (long long)b * lda;
    const unsigned long long* rp = r + (long long)b * ldr;
    // D(i, j): i tokens of a vs j tokens of r.  Diagonal k holds D(i, k - i) for max(0, k-m) <= i <= min(n, k).
    for (int k = 0; k <= n + m; ++k) {
        const int ilo = max(0, k - m), ihi = min(n, k);
        for (int i = ilo + lane; i <= ihi; i += 32) {
            const int j = k - i;
            int v;
            if (i == 0) v = j;
            else if (j == 0) v = i;
            else {
                const int sub = d0[i - 1] + (ap[i - 1] == rp[j - 1] ? 0 : 1);  // D(i-1, j-1) on diagonal k-2
                const int del = d1[i - 1] + 1;                                 // D(i-1, j)   on diagonal k-1
                const int ins = d1[i] + 1;                                     // D(i, j-1)   on diagonal k-1
                v = min(sub, min(del, ins));
            }
            d2[i] = v;
        }
        __syncwarp();
        int* t = d0; d0 = d1; d1 = d2; d2 = t;  // rotate (all lanes agree)
    }
    if (lane == 0) dist[b] = d1[n];  // after the last rotation d1 holds diagonal n + m, whose cell i = n is D(n, m)
}

// floor((len - k) / s) + 1 per conv layer (HF _get_feat_extract_output_lengths with torch.div(rounding_mode="floor")), all
// layers in one launch instead of three element-wise torch kernels per layer on the step's critical path
struct ConvGeom {
    int k[16], s[16], n;
};
__global__ void conv_out_lengths_kernel(const long long* __restrict__ in, int* __restrict__ out, int n, ConvGeom g) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long l = in[i];
    for (int j = 0; j < g.n; ++j) {
        const long long a = l - g.k[j], s = g.s[j];
        long long q = a / s;
        if ((a % s != 0) && ((a < 0) != (s < 0))) --q;   // floor, not truncation, for negative numerators
        l = q + 1;
    }
    out[i] = (int)l;
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

size_t b200asr_normalize_pad_workspace_bytes(int B, long long max_len) {
    if (B <= 0 || max_len <= 0) return 0;
    const long long chunks = (max_len + NORM_CHUNK - 1) / NORM_CHUNK;
    return (size_t)B * (size_t)chunks * 2 * sizeof(double);
}

int b200asr_normalize_pad(const float* packed, const long long* offsets, float* input_values, int* attention_mask,
                          int* frame_lengths, void* workspace, size_t workspace_bytes, int B, int Lmax,
                          const int* conv_kernel, const int* conv_stride, int n_conv, void* stream) {
    B200_REQUIRE(packed && offsets && input_values && workspace, "normalize_pad: null pointer");
    B200_REQUIRE(B > 0 && Lmax > 0, "normalize_pad: bad shape B=%d Lmax=%d", B, Lmax);
    B200_REQUIRE(!frame_lengths || (conv_kernel && conv_stride && n_conv > 0), "normalize_pad: frame lengths need the conv geometry");
    B200_REQUIRE(workspace_bytes >= b200asr_normalize_pad_workspace_bytes(B, Lmax), "normalize_pad: workspace too small");
    const int chunks = (int)(((long long)Lmax + NORM_CHUNK - 1) / NORM_CHUNK);
    cudaStream_t st = as_stream(stream);
    wave_stats_kernel<<<dim3(chunks, B), 256, 0, st>>>(packed, offsets, (double*)workspace, chunks);
    int rc = check_launch("wave_stats");
    if (rc) return rc;
    int gx = (int)(((long long)Lmax + 4095) / 4096);  // 1024 samples per block iteration, 4 iterations per block
    if (gx < 1) gx = 1;
    wave_normalize_pad_kernel<<<dim3(gx, B), 256, 0, st>>>(packed, offsets, (const double*)workspace, input_values,
                                                           attention_mask, frame_lengths, Lmax, chunks, conv_kernel,
                                                           conv_stride, n_conv);
    return check_launch("wave_normalize_pad");
}

int b200asr_words_from_ids(const int* ids, const int* len, int B, int ld, int delim, unsigned long long* words, int* nwords,
                           int ldw, void* stream) {
    B200_REQUIRE(ids && words && nwords, "words_from_ids: null pointer");
    B200_REQUIRE(B > 0 && ld > 0 && ldw > 0, "words_from_ids: bad shape");
    words_from_ids_kernel<<<ceil_div(B, 64), 64, 0, as_stream(stream)>>>(ids, len, B, ld, delim, words, nwords, ldw);
    return check_launch("words_from_ids");
}

int b200asr_edit_distance(const unsigned long long* hyp, const int* hyp_len, const unsigned long long* ref, const int* ref_len,
                          int* dist, int B, int ld_hyp, int ld_ref, void* stream) {
    B200_REQUIRE(hyp && hyp_len && ref && ref_len && dist, "edit_distance: null pointer");
    B200_REQUIRE(B > 0 && ld_hyp > 0 && ld_ref > 0, "edit_distance: bad shape");
    const size_t smem = (size_t)4 * 3 * (ld_hyp + 1) * sizeof(int);
    B200_REQUIRE(smem <= 200 * 1024, "edit_distance: hypothesis rows of %d tokens exceed the shared-memory budget", ld_hyp);
    if (smem > 48 * 1024) cudaFuncSetAttribute(edit_distance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    edit_distance_kernel<<<ceil_div(B, 4), 128, smem, as_stream(stream)>>>(hyp, hyp_len, ref, ref_len, dist, B, ld_hyp, ld_ref);
    return check_launch("edit_distance");
}

int b200asr_conv_out_lengths(const long long* lengths, int* out, int n, const int* conv_kernel, const int* conv_stride,
                             int n_conv, void* stream) {
    B200_REQUIRE(lengths && out && conv_kernel && conv_stride, "conv_out_lengths: null pointer");
    B200_REQUIRE(n >= 1 && n_conv >= 1 && n_conv <= 16, "conv_out_lengths: bad n=%d or n_conv=%d (1..16)", n, n_conv);
    ConvGeom g;
    g.n = n_conv;
    for (int j = 0; j < n_conv; ++j) {
        B200_REQUIRE(conv_stride[j] >= 1 && conv_kernel[j] >= 1, "conv_out_lengths: bad conv geometry at layer %d", j);
        g.k[j] = conv_kernel[j];
        g.s[j] = conv_stride[j];
    }
    conv_out_lengths_kernel<<<(n + 127) / 128, 128, 0, as_stream(stream)>>>(lengths, out, n, g);
    return check_launch("conv_out_lengths");
}

}  // extern "C"
