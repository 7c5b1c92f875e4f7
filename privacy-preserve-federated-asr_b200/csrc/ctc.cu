// CTC loss: fp32 softmax + alpha/beta recursions + gradient w.r.t. the logits, and greedy decode.
// One CTA per (head, utterance).  The frame posteriors of the utterance live in shared memory (T*V*4 bytes);
// warps 0..7 run the alpha recursion while warps 8..15 run beta concurrently (each thread owns SPL consecutive lattice
// states -- one at the sizes of the path -- and the two neighbour states come from the previous thread through shared
// memory, one named barrier per frame and direction).
// The recursions run in the linear domain with a SOFTWARE EXPONENT PER STATE: every state is an fp32 mantissa in [1, 2)
// plus an integer exponent (true value = mantissa * 2^E), renormalised every frame by exact power-of-two scalings -- a
// step is FADD/FMUL and integer ops, no exp/log on the T-step critical path and no warp-wide reduction, the mantissa
// keeps 24 bits whatever the magnitude (torch's log-space fp32 recursion loses bits as |log p| grows) and the dynamic range
// is unbounded.  Logs are taken once, in parallel, when all warps combine alpha*beta into the gradient.
// Latency-bound (T sequential steps).
#include "common.cuh"

namespace b200asr {

namespace {
constexpr int CTC_THREADS = 512;  // lattice kernel: all warps build the posteriors, then 8 warps per recursion
constexpr int LAT_THREADS = 256;  // threads per recursion (alpha: threads 0..255, beta: 256..511)
constexpr float NEG_INF = -INFINITY;
constexpr int E_EMPTY = -(1 << 28);  // exponent of a lane whose states are all zero

// 2^d for d <= 127 (0 below the normal range)
__device__ __forceinline__ float pow2i(int d) { return d < -126 ? 0.f : __int_as_float((d + 127) << 23); }
// m * 2^d for a mantissa m in [1, 2) (or 0) and d <= 0: the exponent field is rewritten by an integer add (no multiply on
// the recursion's dependent chain); below 2^-100 the term is dropped (it is 2^-76 of an fp32 ulp of the sum it joins).
// m == 0 comes with the exponent E_EMPTY: either d < -100, or every operand is empty and d = 0 keeps the zero bits.
__device__ __forceinline__ float scale_pow2(float m, int d) {
    return d < -100 ? 0.f : __int_as_float(__float_as_int(m) + (d << 23));
}
__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// forward (DIR=+1) or backward (DIR=-1) lattice recursion by LAT_THREADS threads (8 warps).
// Thread `tid` owns the SPL consecutive states s = tid*SPL + i; ext[s] = blank (s even) or target[(s-1)/2] (s odd).  The
// states next to a thread's range come from the neighbour thread through a double-buffered shared-memory exchange and ONE
// named barrier per frame (one warp with 8 states per lane, as in round 1 / early round 2, issued ~360 instructions per
// frame from a single warp: 0.7 us per frame, 340 us per utterance; spread over 8 warps a frame is the barrier plus one
// state's ~40 instructions).
// Every state carries its OWN exponent: value = out[t][s] * 2^eout[t][s], mantissa in [1, 2) (or 0 with E_EMPTY).
// (Round 1 shared one exponent per lane: with peaked posteriors -- emission ratios of 2^-60 per frame -- two states of a
// lane drift more than the 2^-126 fp32 range apart within a few frames and the smaller one, which may carry the whole
// posterior mass once multiplied with the opposite recursion, was flushed to zero: wrong gradients at full-size configs.)
template <int SPL, int DIR>
__device__ void lattice_pass(const float* __restrict__ pr, int V, const int* __restrict__ tgt, int L, int Tb, int blank,
                             float* __restrict__ out /* [T][Sld] */, int Sld, int* __restrict__ eout /* [T][Sld] */,
                             int2* __restrict__ xb /* [2][LAT_THREADS][2] */, int tid, int bar_id) {
    // Sld = LAT_THREADS * SPL here; state s = tid * SPL + i lives at [t][i * LAT_THREADS + tid]
    constexpr int NT = LAT_THREADS;
    const int S = 2 * L + 1;
    int lab[SPL];
    bool skip[SPL];  // may take the 2-step transition (from s-2 forward / from s+2 backward)
#pragma unroll
    for (int i = 0; i < SPL; ++i) {
        const int s = tid * SPL + i;
        lab[i] = blank;
        skip[i] = false;
        if (s < S && (s & 1)) {
            const int l = tgt[(s - 1) >> 1];
            lab[i] = l;
            if (DIR > 0) skip[i] = (s >= 3) && (tgt[(s - 3) >> 1] != l);
            else skip[i] = (s + 2 < S) && (tgt[(s + 1) >> 1] != l);
        }
    }
    float a[SPL];
    int E[SPL];
    auto normalise = [](float v, int base, float& m, int& e) {   // v >= 0 -> mantissa in [1, 2) and exponent (base + k)
        const int bits = __float_as_int(v);
        const bool live = v > 0.f;   // (v is a sum of [1, 2) mantissas times a [1, 2) emission: normal or exactly 0)
        m = live ? __int_as_float((bits & 0x007fffff) | 0x3f800000) : 0.f;
        e = live ? base + (bits >> 23) - 127 : E_EMPTY;
    };
    // the two states a neighbour needs: slot 0 = the state next to its range, slot 1 = the one after (SPL == 1: the
    // neighbour's neighbour publishes that one itself)
    auto publish = [&](int buf) {
        int2* x = xb + (buf * NT + tid) * 2;
        if (DIR > 0) {
            x[0] = make_int2(__float_as_int(a[SPL - 1]), E[SPL - 1]);
            if (SPL > 1) x[1] = make_int2(__float_as_int(a[SPL > 1 ? SPL - 2 : 0]), E[SPL > 1 ? SPL - 2 : 0]);
        } else {
            x[0] = make_int2(__float_as_int(a[0]), E[0]);
            if (SPL > 1) x[1] = make_int2(__float_as_int(a[SPL > 1 ? 1 : 0]), E[SPL > 1 ? 1 : 0]);
        }
    };
    const int t0 = DIR > 0 ? 0 : Tb - 1;
#pragma unroll
    for (int i = 0; i < SPL; ++i) {
        const int s = tid * SPL + i;
        bool init;
        if (DIR > 0) init = (s == 0) || (s == 1 && S > 1);
        else init = (s == S - 1) || (s == S - 2 && S > 1);
        if (init && s < S) {   // pr holds log2-probabilities: mantissa 2^frac in [1, 2), exponent floor(l2)
            const float l2 = pr[t0 * V + lab[i]];
            const float fl = fmaxf(floorf(l2), -1.0e8f);   // (-inf logits: probability 0)
            a[i] = ex2_fast(l2 - fl);
            E[i] = a[i] > 0.f ? (int)fl : E_EMPTY;
        } else {
            a[i] = 0.f;
            E[i] = E_EMPTY;
        }
        if (s < S) {   // thread-contiguous layout [t][i][tid]: every store of a warp is one 128-byte line
            out[(long long)t0 * Sld + i * NT + tid] = a[i];
            eout[(long long)t0 * Sld + i * NT + tid] = E[i];
        }
    }
    publish(0);
    const int nb = DIR > 0 ? tid - 1 : tid + 1;            // the neighbour thread whose states precede this range
    const int nb2 = SPL > 1 ? nb : (DIR > 0 ? tid - 2 : tid + 2);
    const bool has1 = nb >= 0 && nb < NT, has2 = nb2 >= 0 && nb2 < NT;
    for (int step = 1; step < Tb; ++step) {
        const int t = DIR > 0 ? step : Tb - 1 - step;
        float e[SPL];  // emission probabilities of this frame (mantissa, exponent): independent of the recursion, issued first
        int ee[SPL];   // (log2-domain storage: no posterior underflows, whatever the logit range -- as torch's log-space lattice)
#pragma unroll
        for (int i = 0; i < SPL; ++i) {
            const float l2 = pr[t * V + lab[i]];
            const float fl = fmaxf(floorf(l2), -1.0e8f);   // (-inf logits: probability 0)
            e[i] = ex2_fast(l2 - fl);
            ee[i] = (int)fl;
        }
        // the previous frame's states of every thread of this direction are published
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(NT) : "memory");
        const int2* xp = xb + ((step - 1) & 1) * NT * 2;
        int2 x1 = make_int2(0, E_EMPTY), x2 = make_int2(0, E_EMPTY);
        if (has1) x1 = xp[nb * 2];
        if (has2) x2 = xp[nb2 * 2 + (SPL > 1 ? 1 : 0)];
        const float n1 = __int_as_float(x1.x), n2 = __int_as_float(x2.x);
        const int En1 = x1.y, En2 = x2.y;
        float ao[SPL];
        int Eo[SPL];
#pragma unroll
        for (int i = 0; i < SPL; ++i) { ao[i] = a[i]; Eo[i] = E[i]; }
#pragma unroll
        for (int i = 0; i < SPL; ++i) {
            float m1, m2;
            int e1, e2;
            if (DIR > 0) {
                if (i >= 1) { m1 = ao[i - 1]; e1 = Eo[i - 1]; } else { m1 = n1; e1 = En1; }
                if (i >= 2) { m2 = ao[i - 2]; e2 = Eo[i - 2]; } else if (i == 1) { m2 = n1; e2 = En1; } else { m2 = n2; e2 = En2; }
            } else {
                if (i + 1 < SPL) { m1 = ao[i + 1]; e1 = Eo[i + 1]; } else { m1 = n1; e1 = En1; }
                if (i + 2 < SPL) { m2 = ao[i + 2]; e2 = Eo[i + 2]; } else if (i + 2 == SPL) { m2 = n1; e2 = En1; } else { m2 = n2; e2 = En2; }
            }
            if (!skip[i]) { m2 = 0.f; e2 = E_EMPTY; }
            const int s = tid * SPL + i;
            const int Ec = max(Eo[i], max(e1, e2));   // common exponent of this state's operands
            // (an empty operand has mantissa 0 and E_EMPTY - Ec underflows to a factor 0; all empty: 0 * 1 = 0)
            const float v = (scale_pow2(ao[i], Eo[i] - Ec) + scale_pow2(m1, e1 - Ec) + scale_pow2(m2, e2 - Ec)) * e[i];
            normalise(s < S ? v : 0.f, Ec + ee[i], a[i], E[i]);
        }
        publish(step & 1);
#pragma unroll
        for (int i = 0; i < SPL; ++i) {
            if (tid * SPL + i < S) {
                out[(long long)t * Sld + i * NT + tid] = a[i];
                eout[(long long)t * Sld + i * NT + tid] = E[i];
            }
        }
    }
}

template <int DIR>
__device__ void lattice_dispatch(int spl, const float* pr, int V, const int* tgt, int L, int Tb, int blank, float* out,
                                 int Sld, int* eout, int2* xb, int tid, int bar_id) {
    switch (spl) {
        case 1: lattice_pass<1, DIR>(pr, V, tgt, L, Tb, blank, out, Sld, eout, xb, tid, bar_id); break;
        case 2: lattice_pass<2, DIR>(pr, V, tgt, L, Tb, blank, out, Sld, eout, xb, tid, bar_id); break;
        default: lattice_pass<4, DIR>(pr, V, tgt, L, Tb, blank, out, Sld, eout, xb, tid, bar_id); break;
    }
}
}  // namespace

// fp32 softmax of one frame by one warp (same operation order wherever it is used: the lattice kernel and the gradient
// kernel must see bit-identical posteriors).  Up to 4 values per lane stay in registers: one global read of the row.
template <typename T, typename F>
__device__ __forceinline__ void frame_softmax(const T* __restrict__ xr, int V, int lane, F&& emit) {
    float xv[4];
    const bool cached = V <= 128;
    float mx = NEG_INF;
    if (cached) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int v = lane + 32 * k;
            xv[k] = v < V ? to_f<T>(xr[v]) : NEG_INF;
            mx = fmaxf(mx, xv[k]);
        }
    } else {
        for (int v = lane; v < V; v += 32) mx = fmaxf(mx, to_f<T>(xr[v]));
    }
    mx = warp_max(mx);
    float sum = 0.f;
    if (cached) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (lane + 32 * k < V) sum += expf(xv[k] - mx);
    } else {
        for (int v = lane; v < V; v += 32) sum += expf(to_f<T>(xr[v]) - mx);
    }
    sum = warp_sum(sum);
    const float lz = mx + logf(sum);
    constexpr float LOG2E_ = 1.4426950408889634f;
    if (cached) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (lane + 32 * k < V) emit(lane + 32 * k, expf(xv[k] - lz), (xv[k] - lz) * LOG2E_);
    } else {
        for (int v = lane; v < V; v += 32) emit(v, expf(to_f<T>(xr[v]) - lz), (to_f<T>(xr[v]) - lz) * LOG2E_);
    }
}

// workspace per (head, utterance), in floats: alpha, beta mantissas [T][Sld], their per-state exponents [T][Sld] int32 each,
// then 8 floats of scalars (0: the raw negative log-likelihood for the gradient kernel)
__host__ __device__ inline long long ctc_ws_stride(int T_, int Sld) { return 4LL * T_ * Sld + 8; }

// Lattice kernel: one CTA per (head, utterance).  Phase 1: posteriors of the utterance into shared memory (all warps);
// phase 2: alpha (warps 0..7) and beta (warps 8..15) concurrently; then the loss.  The gradient is a separate, frame-parallel
// kernel: inside this CTA it took 350 of 510 us (8 utterances = 8 busy SMs) at the sizes of the path.
template <typename T>
__global__ void __launch_bounds__(CTC_THREADS) ctc_kernel(const T* __restrict__ logits, const int* __restrict__ targets,
                                                          const int* __restrict__ tgt_off, const int* __restrict__ tgt_len,
                                                          const int* __restrict__ in_len, const float* __restrict__ grad_scale,
                                                          float* __restrict__ loss, T* __restrict__ grad,
                                                          float* __restrict__ ws, int B, int T_, int V, int max_tgt,
                                                          int blank, int zero_infinity, int spl) {
    extern __shared__ float sm[];
    const int b = blockIdx.x, head = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int Sld = LAT_THREADS * spl;   // states per frame in the workspace: state s at [(s % spl) * LAT_THREADS + s / spl]
    const int L = min(tgt_len[b], max_tgt);
    const int S = 2 * L + 1;
    const int Tb = max(0, min(in_len[b], T_));
    int2* xb = reinterpret_cast<int2*>(sm);          // [2 directions][2 buffers][LAT_THREADS][2] boundary-state exchange
    float* pr = sm + 2 * 2 * LAT_THREADS * 2 * 2;    // [T][V] log2 of the softmax probabilities
    int* tgt_s = (int*)(pr + (size_t)T_ * V);        // [max_tgt + 1]
    const long long hb = (long long)head * B + b;
    const T* x = logits + hb * T_ * V;
    float* alpha = ws + hb * ctc_ws_stride(T_, Sld);
    float* beta = alpha + (long long)T_ * Sld;
    int* ea = reinterpret_cast<int*>(beta + (long long)T_ * Sld);
    int* eb = ea + (long long)T_ * Sld;

    for (int i = threadIdx.x; i < L; i += blockDim.x) tgt_s[i] = targets[tgt_off[b] + i];
    // ---- phase 1: fp32 softmax of the valid frames
    for (int t = warp; t < Tb; t += nwarps)
        frame_softmax<T>(x + (long long)t * V, V, lane, [&](int v, float, float l2v) { pr[t * V + v] = l2v; });
    __syncthreads();
    // ---- phase 2: alpha (warps 0..7) and beta (warps 8..15) concurrently, each with its own named barrier
    if (Tb > 0) {
        if (threadIdx.x < LAT_THREADS)
            lattice_dispatch<+1>(spl, pr, V, tgt_s, L, Tb, blank, alpha, Sld, ea, xb, threadIdx.x, 1);
        else
            lattice_dispatch<-1>(spl, pr, V, tgt_s, L, Tb, blank, beta, Sld, eb, xb + 2 * LAT_THREADS * 2,
                                 threadIdx.x - LAT_THREADS, 2);
    }
    __threadfence_block();
    __syncthreads();
    // ---- phase 3: loss and gradient
    constexpr float LN2 = 0.6931471805599453f;
    float nll = INFINITY;
    float lp_frac = 0.f;   // log2 P(y) = lp_frac + lp_exp, kept split: the gradient kernel subtracts the integer parts exactly
    int lp_exp = 0;
    if (Tb > 0) {
        const long long r = (long long)(Tb - 1);
        auto at = [&](int st) { return r * Sld + (st % spl) * LAT_THREADS + st / spl; };
        const float a1 = alpha[at(S - 1)];
        const int e1 = ea[at(S - 1)];
        const float a2 = S > 1 ? alpha[at(S - 2)] : 0.f;
        const int e2 = S > 1 ? ea[at(S - 2)] : E_EMPTY;
        const int em = max(e1, e2);
        const float tot = a1 * pow2i(e1 - em) + a2 * pow2i(e2 - em);
        if (tot > 0.f) {
            nll = -(logf(tot) + (float)em * LN2);
            lp_frac = log2f(tot);
            lp_exp = em;
        }
    }
    if (Tb == 0 && L == 0) nll = 0.f;  // empty input, empty target (torch convention)
    const bool infeasible = !(nll < INFINITY);
    if (threadIdx.x == 0) {
        loss[hb] = (infeasible && zero_infinity) ? 0.f : nll;
        float* sc = reinterpret_cast<float*>(eb + (long long)T_ * Sld);   // scalars for the gradient kernel
        sc[0] = nll;
        sc[1] = lp_frac;
        reinterpret_cast<int*>(sc)[2] = lp_exp;
    }
}

// Gradient kernel: one warp per frame of one (head, utterance); reads alpha / beta of that frame, recomputes the frame's
// posteriors (bit-identical to the lattice kernel's), accumulates the label occupancies in 2^-40 fixed point (integer
// atomics commute: reproducible) and writes d nll / d logits.
template <typename T>
__global__ void __launch_bounds__(256) ctc_grad_kernel(const T* __restrict__ logits, const int* __restrict__ targets,
                                                       const int* __restrict__ tgt_off, const int* __restrict__ tgt_len,
                                                       const int* __restrict__ in_len, const float* __restrict__ grad_scale,
                                                       T* __restrict__ grad, const float* __restrict__ ws, int n_heads,
                                                       int B, int T_, int V, int max_tgt, int blank, int zero_infinity,
                                                       int spl) {
    extern __shared__ unsigned long long acc_all[];   // [8 warps][V] occupancy sums + [8][2][V] floats: posteriors and their log2
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long frame = (long long)blockIdx.x * 8 + warp;
    if (frame >= (long long)n_heads * B * T_) return;
    const long long hb = frame / T_;
    const int t = (int)(frame - hb * T_);
    const int b = (int)(hb % B);
    const int Sld = LAT_THREADS * spl;
    const int L = min(tgt_len[b], max_tgt);
    const int S = 2 * L + 1;
    const int Tb = max(0, min(in_len[b], T_));
    const float* alpha = ws + hb * ctc_ws_stride(T_, Sld);
    const float* beta = alpha + (long long)T_ * Sld;
    const int* ea = reinterpret_cast<const int*>(beta + (long long)T_ * Sld);
    const int* eb = ea + (long long)T_ * Sld;
    const float* scal = reinterpret_cast<const float*>(eb + (long long)T_ * Sld);
    const float nll = scal[0], lp_frac = scal[1];
    const int lp_exp = reinterpret_cast<const int*>(scal)[2];
    const bool infeasible = !(nll < INFINITY);
    constexpr float LOG2E = 1.4426950408889634f;
    T* g = grad + hb * T_ * V;
    if (t >= Tb || infeasible) {
        // torch: gradient of padded frames is 0; zero_infinity zeroes infeasible utterances (NaN otherwise)
        for (int v = lane; v < V; v += 32) g[(long long)t * V + v] = from_f<T>((t < Tb && !zero_infinity) ? NAN : 0.f);
        return;
    }
    unsigned long long* acc = acc_all + warp * V;
    float* pr = reinterpret_cast<float*>(acc_all + 8 * V) + warp * 2 * V;   // [V] probabilities, [V] their log2
    float* pl2 = pr + V;
    const int* tgt = targets + tgt_off[b];
    const float gs = grad_scale ? grad_scale[hb] : 1.f;
    frame_softmax<T>(logits + (hb * T_ + t) * V, V, lane, [&](int v, float pv, float l2v) { pr[v] = pv; pl2[v] = l2v; });
    const float* ar = alpha + (long long)t * Sld;
    const float* br = beta + (long long)t * Sld;
    const int* ear = ea + (long long)t * Sld;
    const int* ebr = eb + (long long)t * Sld;
    __syncwarp();   // pl2[] of the other lanes' labels is read below
    // q(s) = log2( alpha_t(s) beta_t(s) / (p_t(label_s) P(y)) ) = log2 of the state posterior: mantissa product + the two state
    // exponents - log2 p - log2 P (integer parts subtracted exactly: the result is a small number, whatever |log P| is).  The division by p comes FIRST: with peaked posteriors a state whose label has p = 2^-60 carries
    // alpha beta = 2^-60 of the frame's largest product and still holds the whole posterior mass of its label (a fixed-point
    // sum relative to max alpha beta, as in round 1, drops exactly those states).
    float mx = NEG_INF;
    for (int s = lane; s < S; s += 32) {
        const int ix = (s % spl) * LAT_THREADS + s / spl;
        const float ab = ar[ix] * br[ix];
        if (ab > 0.f) {
            const int l = (s & 1) ? tgt[(s - 1) >> 1] : blank;
            mx = fmaxf(mx, (log2f(ab) - pl2[l] - lp_frac) + (float)(ear[ix] + ebr[ix] - lp_exp));
        }
    }
    mx = warp_max(mx);
    for (int v = lane; v < V; v += 32) acc[v] = 0ull;
    __syncwarp();
    for (int s = lane; s < S; s += 32) {
        const int ix = (s % spl) * LAT_THREADS + s / spl;
        const float ab = ar[ix] * br[ix];
        if (ab > 0.f) {
            const int l = (s & 1) ? tgt[(s - 1) >> 1] : blank;
            const float rel2 = (log2f(ab) - pl2[l] - lp_frac) + (float)(ear[ix] + ebr[ix] - lp_exp) - mx;  // <= 0
            atomicAdd(&acc[l], (unsigned long long)(exp2f(rel2) * 1099511627776.0f));  // in [0, 2^40]: integer sums commute
        }
    }
    __syncwarp();
    // d nll / d logit = softmax - occupancy,  occupancy_t(v) = sum_{s: lab = v} alpha_t(s) beta_t(s) / (P p_t(v)),  P = exp(-nll)
    // (ATen ctc_loss backward form)
    const float fac = exp2f(mx - 40.0f);   // (mx = log2 of the frame's largest state posterior: <= 0 up to rounding)
    for (int v = lane; v < V; v += 32) {
        const float pv = pr[v];
        const float occ = (float)acc[v] * fac;
        g[(long long)t * V + v] = from_f<T>((pv - occ) * gs);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) greedy_kernel(const T* __restrict__ logits, const int* __restrict__ in_len,
                                                     int* __restrict__ ids, int* __restrict__ collapsed,
                                                     int* __restrict__ clen, int T_, int V, int blank) {
    const int b = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int Tb = in_len ? min(in_len[b], T_) : T_;
    const T* x = logits + (long long)b * T_ * V;
    for (int t = warp; t < T_; t += nwarps) {
        float best = NEG_INF;
        int bi = 0x7fffffff;
        for (int v = lane; v < V; v += 32) {
            const float f = to_f<T>(x[(long long)t * V + v]);
            if (f > best || (f == best && v < bi)) { best = f; bi = v; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (lane == 0) ids[(long long)b * T_ + t] = bi;
    }
    __threadfence_block();
    __syncthreads();
    if (threadIdx.x == 0) {
        int n = 0, prev = -1;
        for (int t = 0; t < Tb; ++t) {
            const int id = ids[(long long)b * T_ + t];
            if (id != prev && id != blank) collapsed[(long long)b * T_ + n++] = id;
            prev = id;
        }
        clen[b] = n;
        for (int t = n; t < T_; ++t) collapsed[(long long)b * T_ + t] = -1;
    }
}

}  // namespace b200asr

using namespace b200asr;

extern "C" {

static int ctc_spl(int max_tgt_len) {   // lattice states per thread (1, 2, 4; > 4: unsupported)
    const int S = 2 * (max_tgt_len < 1 ? 1 : max_tgt_len) + 1;
    int spl = 1;
    while (spl * LAT_THREADS < S) spl *= 2;
    return spl;
}

size_t b200asr_ctc_workspace_bytes(int n_heads, int B, int T, int max_tgt_len) {
    return (size_t)n_heads * B * (size_t)ctc_ws_stride(T, LAT_THREADS * ctc_spl(max_tgt_len)) * sizeof(float);
}

int b200asr_ctc_fwd_bwd(const void* logits, const int* targets, const int* tgt_off, const int* tgt_len,
                        const int* in_len, const float* grad_scale, float* loss, void* grad, void* workspace,
                        size_t workspace_bytes, int n_heads, int B, int T_, int V, int max_tgt_len, int blank,
                        int zero_infinity, int dtype, void* stream) {
    B200_REQUIRE(logits && targets && tgt_off && tgt_len && in_len && loss && workspace, "ctc: null pointer");
    B200_REQUIRE(n_heads >= 1 && B >= 1 && T_ >= 1 && V >= 2, "ctc: bad shape");
    B200_REQUIRE(blank >= 0 && blank < V, "ctc: blank out of range");
    if (max_tgt_len < 1) max_tgt_len = 1;
    B200_REQUIRE(workspace_bytes >= b200asr_ctc_workspace_bytes(n_heads, B, T_, max_tgt_len), "ctc: workspace too small");
    const int spl = ctc_spl(max_tgt_len);
    if (spl > 4) {
        set_error("ctc: target length %d exceeds the supported 511", max_tgt_len);
        return B200ASR_ERR_UNSUPPORTED;
    }
    const size_t smem = ((size_t)T_ * V + (max_tgt_len + 1) + 2 * 2 * LAT_THREADS * 2 * 2) * sizeof(float) + 8;
    if (smem > 227 * 1024) {
        set_error("ctc: T*V=%d*%d log-probabilities do not fit in shared memory", T_, V);
        return B200ASR_ERR_UNSUPPORTED;
    }
    dim3 grid(B, n_heads);
    cudaStream_t st = as_stream(stream);
    B200_DISPATCH_DTYPE(dtype, {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(ctc_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) {
                set_error("ctc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
                return B200ASR_ERR_CUDA;
            }
        }
        ctc_kernel<T><<<grid, CTC_THREADS, smem, st>>>((const T*)logits, targets, tgt_off, tgt_len, in_len, grad_scale, loss,
                                                        (T*)grad, (float*)workspace, B, T_, V, max_tgt_len, blank,
                                                        zero_infinity, spl);
        if (grad) {
            const long long frames = (long long)n_heads * B * T_;
            const size_t gsm = (size_t)8 * V * (sizeof(unsigned long long) + 2 * sizeof(float));
            if (gsm > 48 * 1024) {
                set_error("ctc: vocabulary %d too large for the gradient kernel", V);
                return B200ASR_ERR_UNSUPPORTED;
            }
            ctc_grad_kernel<T><<<(unsigned)((frames + 7) / 8), 256, gsm, st>>>((const T*)logits, targets, tgt_off, tgt_len, in_len,
                                                                            grad_scale, (T*)grad, (const float*)workspace,
                                                                            n_heads, B, T_, V, max_tgt_len, blank,
                                                                            zero_infinity, spl);
        }
    });
    return check_launch("ctc_fwd_bwd");
}

int b200asr_greedy_decode(const void* logits, const int* in_len, int* out_ids, int* collapsed, int* collapsed_len, int B,
                          int T_, int V, int blank, int dtype, void* stream) {
    B200_REQUIRE(logits && out_ids && collapsed && collapsed_len, "greedy_decode: null pointer");
    B200_REQUIRE(B >= 1 && T_ >= 1 && V >= 1, "greedy_decode: bad shape");
    B200_DISPATCH_DTYPE(dtype, {
        greedy_kernel<T><<<B, 256, 0, as_stream(stream)>>>((const T*)logits, in_len, out_ids, collapsed, collapsed_len, T_, V,
                                                          blank);
    });
    return check_launch("greedy_decode");
}

}  // extern "C"
