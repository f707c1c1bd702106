// Network kernels of the learner step, exact-precision (fp32 FFMA) path:
// conv torso fwd/bwd (model.py:13-21,44-46), LSTM unroll fwd + BPTT (model.py:52-60), dueling
// head fwd/bwd (model.py:26-36,62-64), n-step double-Q target / TD / priority
// (learner.py:86-102), centered RMSprop (learner.py:51,100).
#include "gemm.cuh"

#include <mutex>

namespace r2d2 {

// ======================================================================== small helpers
__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.f / (1.f + expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
    return v;
}

// deterministic block reduction (blockDim.x <= 1024); result valid in thread 0
__device__ float block_sum(float v, float* red /*[32]*/) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        v = lane < (blockDim.x + 31) / 32 ? red[lane] : 0.f;
        v = warp_sum(v);
    }
    return v;
}

// out[i] = sum_z part[z*slab + i]   (split-K partial reduction, fixed order)
__global__ void reduce_partials_kernel(const float* __restrict__ part, float* __restrict__ out,
                                       int64_t n, int splits) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += part[(int64_t)z * n + i];
        out[i] = s;
    }
}

// db[n] = sum_m dy[m*N + n]: 8 columns (one 32-byte sector per row) per CTA, 128 row lanes
// stride over the rows with four independent accumulators, then a fixed-order reduction
__global__ void __launch_bounds__(1024)
colsum_kernel(const float* __restrict__ dy, float* __restrict__ db, int M, int N,
              float* __restrict__ db_copy = nullptr) {
    // 32 columns x 32 row lanes per block: a warp reads 128 contiguous bytes of one row
    __shared__ float red[32][33];
    const int col = threadIdx.x & 31, rl = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + col;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (n < N) {
        const float* p = dy + n;
        int m = rl;
        for (; m + 96 < M; m += 128) {
            s0 += p[(int64_t)m * N];
            s1 += p[(int64_t)(m + 32) * N];
            s2 += p[(int64_t)(m + 64) * N];
            s3 += p[(int64_t)(m + 96) * N];
        }
        for (; m < M; m += 32) s0 += p[(int64_t)m * N];
    }
    red[rl][col] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    for (int h = 16; h >= 1; h >>= 1) {
        if (rl < h) red[rl][col] += red[rl + h][col];
        __syncthreads();
    }
    if (rl == 0 && n < N) {
        db[n] = red[0][col];
        if (db_copy) db_copy[n] = red[0][col];   // e.g. bias_ih and bias_hh share one gradient
    }
}

// per-channel sum over frames and pixels of an NCHW gradient: db[c] = sum_{r,p} dz[r,c,p], C = 32.
// CTA z walks frames z, z+Z, ...; thread t keeps the float4s t, t+256, ... of a frame in
// registers (the same channel/pixel every frame), so the frame loop is pure coalesced 16-byte
// loads + adds.  The register image is then laid out in shared memory and each warp sums four
// channels in a fixed order -> part[z*32 + c]; chansum_final adds the Z partials in fixed order
// (deterministic, no atomics).
constexpr int CS_Z = R2D2_CHANSUM_PART_FLOATS / 32;   // 296 CTAs = 2 per SM
template <int PIX>
__global__ void __launch_bounds__(256)
chansum_partial_kernel(const float* __restrict__ dz, float* __restrict__ part, int R) {
    constexpr int Q = 32 * PIX / 4;              // float4s per frame
    constexpr int NQ = (Q + 255) / 256;
    extern __shared__ float4 img[];              // [Q]
    float4 acc[NQ];
#pragma unroll
    for (int j = 0; j < NQ; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = blockIdx.x; r < R; r += gridDim.x) {
        const float4* f = reinterpret_cast<const float4*>(dz + (int64_t)r * 32 * PIX);
#pragma unroll
        for (int j = 0; j < NQ; ++j) {
            const int q = threadIdx.x + 256 * j;
            if (q < Q) {
                const float4 v = __ldg(f + q);
                acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NQ; ++j) {
        const int q = threadIdx.x + 256 * j;
        if (q < Q) img[q] = acc[j];
    }
    __syncthreads();
    const float* flat = reinterpret_cast<const float*>(img);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int c = w; c < 32; c += 8) {
        float s = 0.f;
        for (int i = lane; i < PIX; i += 32) s += flat[c * PIX + i];
        s = warp_sum(s);
        if (lane == 0) part[blockIdx.x * 32 + c] = s;
    }
}
// one warp per channel: lanes stride over the Z partials, fixed-order shuffle tree
__global__ void __launch_bounds__(1024)
chansum_final_kernel(const float* __restrict__ part, float* __restrict__ db, int Z) {
    const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;
    float s = 0.f;
    for (int z = lane; z < Z; z += 32) s += part[z * 32 + c];
    s = warp_sum(s);
    if (lane == 0) db[c] = s;
}
template <int PIX>
static int chansum_launch(const float* dz, float* db, int R, float* part, cudaStream_t st) {
    constexpr int smem = 32 * PIX * 4;
    static bool attr_set = false;
    if (!attr_set) {
        R2D2_CUDA_OK(cudaFuncSetAttribute(chansum_partial_kernel<PIX>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    const int Z = R < CS_Z ? R : CS_Z;
    R2D2_LAUNCH("chansum_partial_kernel", st, chansum_partial_kernel<PIX><<<Z, 256, smem, st>>>(dz, part, R));
    R2D2_LAUNCH("chansum_final_kernel", st, chansum_final_kernel<<<1, 1024, 0, st>>>(part, db, Z));
    return R2D2_OK;
}
static int chansum_dispatch(const float* dz, float* db, int R, int PIX, float* part, cudaStream_t st) {
    switch (PIX) {
        case 49: return chansum_launch<49>(dz, db, R, part, st);
        case 81: return chansum_launch<81>(dz, db, R, part, st);
        case 400: return chansum_launch<400>(dz, db, R, part, st);
    }
    return r2d2::set_err(R2D2_EINVAL, "chansum: PIX must be 49, 81 or 400 (the three conv layers), got %d", PIX);
}

// d[i] = a[i] > 0 ? d[i] : 0
__global__ void relu_mask_kernel(float* __restrict__ d, const float* __restrict__ a, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        if (!(a[i] > 0.f)) d[i] = 0.f;
}

// ======================================================================== conv data-gradient
// dIn[r,ci,y,x] = relu'(act[r,ci,y,x]) * sum_{co,ky,kx} dZ[r,co,(y-ky)/S,(x-kx)/S] W[co,ci,ky,kx]
// One CTA per frame; dZ and the (co,ky,kx)-major / ci-minor weights are staged in smem.
template <int KH, int HIN, int HOUT, int STRIDE>
__global__ void __launch_bounds__(256)
conv_dgrad_kernel(const float* __restrict__ dz, const float* __restrict__ w,
                  const float* __restrict__ act, float* __restrict__ din) {
    constexpr int C = 32, PO = HOUT * HOUT, PI = HIN * HIN, TAPS = KH / STRIDE;
    extern __shared__ __align__(16) float sm[];
    float* s_w = sm;                      // [(co*KH+ky)*KH+kx][ci]
    float* s_dz = sm + C * KH * KH * C;   // [co][PO]
    const int r = blockIdx.x;
    for (int i = threadIdx.x; i < C * C * KH * KH; i += blockDim.x) {
        // source index i = ((co*C + ci)*KH + ky)*KH + kx
        const int kx = i % KH, ky = (i / KH) % KH, ci = (i / (KH * KH)) % C, co = i / (KH * KH * C);
        s_w[((co * KH + ky) * KH + kx) * C + ci] = __ldg(w + i);
    }
    for (int i = threadIdx.x; i < C * PO; i += blockDim.x) s_dz[i] = dz[(int64_t)r * C * PO + i];
    __syncthreads();
    for (int item = threadIdx.x; item < PI * 4; item += blockDim.x) {
        const int pix = item >> 2, cg = (item & 3) * 8;
        const int y = pix / HIN, x = pix - y * HIN;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
        for (int jy = 0; jy < TAPS; ++jy) {
            const int ky = (y % STRIDE) + jy * STRIDE;
            const int oy = (y - ky) / STRIDE;
            if (y - ky < 0 || oy >= HOUT) continue;
#pragma unroll
            for (int jx = 0; jx < TAPS; ++jx) {
                const int kx = (x % STRIDE) + jx * STRIDE;
                const int ox = (x - kx) / STRIDE;
                if (x - kx < 0 || ox >= HOUT) continue;
                for (int co = 0; co < C; ++co) {
                    const float g = s_dz[co * PO + oy * HOUT + ox];
                    const float4* wp =
                        reinterpret_cast<const float4*>(s_w + ((co * KH + ky) * KH + kx) * C + cg);
                    const float4 w0 = wp[0], w1 = wp[1];
                    acc[0] = fmaf(g, w0.x, acc[0]); acc[1] = fmaf(g, w0.y, acc[1]);
                    acc[2] = fmaf(g, w0.z, acc[2]); acc[3] = fmaf(g, w0.w, acc[3]);
                    acc[4] = fmaf(g, w1.x, acc[4]); acc[5] = fmaf(g, w1.y, acc[5]);
                    acc[6] = fmaf(g, w1.z, acc[6]); acc[7] = fmaf(g, w1.w, acc[7]);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int64_t o = ((int64_t)r * C + cg + j) * PI + pix;
            din[o] = act[o] > 0.f ? acc[j] : 0.f;
        }
    }
}

// ======================================================================== LSTM
// One step of the recurrence for UPB hidden units x all four gates x a 64-row batch tile.
//   pre = xproj_t[b, g*H+u] + sum_k h_prev[b,k] * W_hh[g*H+u, k];  i,f,o = sigmoid, g = tanh
//   c = f*c_prev + i*g;  h = o*tanh(c)                              (model.py:55)
// 256 threads: 4-way split of every 64-wide K chunk x (16 row groups x 4 column groups) of
// 4x4 register tiles; partials are summed through smem before the pointwise update.
constexpr int LSTM_UPB = 4, LSTM_BT = 64, LSTM_KC = 64;

__global__ void __launch_bounds__(256)
lstm_step_fwd_kernel(const float* __restrict__ xproj_t, const float* __restrict__ w_hh,
                     const float* __restrict__ h_prev, const float* __restrict__ c_prev,
                     float* __restrict__ h_out, float* __restrict__ c_out,
                     float* __restrict__ gates_out, int B, int H) {
    constexpr int NC = 4 * LSTM_UPB;  // 16 gate columns
    __shared__ __align__(16) float Hs[LSTM_KC][LSTM_BT + 4];
    __shared__ __align__(16) float Ws[LSTM_KC][NC + 4];
    const int tid = threadIdx.x;
    const int u0 = blockIdx.x * LSTM_UPB, b0 = blockIdx.y * LSTM_BT;
    const int kq = tid >> 6, tb = (tid & 63) >> 2, tc = tid & 3;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int kc = 0; kc < H; kc += LSTM_KC) {
        {   // stage h_prev[b0.., kc..] as Hs[k][b] and the 16 weight rows as Ws[k][c]
            const int k = tid & 63;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int b = (tid >> 6) + 4 * j;
                Hs[k][b] = (b0 + b < B && kc + k < H) ? h_prev[(int64_t)(b0 + b) * H + kc + k] : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = (tid >> 6) + 4 * j;  // c = gate*UPB + unit
                const int row = (c / LSTM_UPB) * H + u0 + (c % LSTM_UPB);
                Ws[k][c] = (kc + k < H) ? __ldg(w_hh + (int64_t)row * H + kc + k) : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < LSTM_KC / 4; ++kk) {
            const int k = kq * (LSTM_KC / 4) + kk;
            const float4 hv = *reinterpret_cast<const float4*>(&Hs[k][tb * 4]);
            const float4 wv = *reinterpret_cast<const float4*>(&Ws[k][tc * 4]);
            const float a[4] = {hv.x, hv.y, hv.z, hv.w}, w[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
        }
        __syncthreads();
    }
    // cross-split reduction through smem: red[kq][b][c]
    float* red = &Hs[0][0];  // 64*68 floats >= 4*64*16
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) red[(kq * LSTM_BT + tb * 4 + i) * NC + tc * 4 + j] = acc[i][j];
    __syncthreads();
    const int b = tid >> 2, j = tid & 3;  // one (row, unit) per thread
    if (b0 + b >= B) return;
    float pre[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int c = g * LSTM_UPB + j;
        float s = red[(0 * LSTM_BT + b) * NC + c] + red[(1 * LSTM_BT + b) * NC + c];
        s += red[(2 * LSTM_BT + b) * NC + c] + red[(3 * LSTM_BT + b) * NC + c];
        pre[g] = s + xproj_t[(int64_t)(b0 + b) * 4 * H + g * H + u0 + j];
    }
    const float ig = sigmoidf_acc(pre[0]), fg = sigmoidf_acc(pre[1]);
    const float gg = tanhf(pre[2]), og = sigmoidf_acc(pre[3]);
    const int64_t o = (int64_t)(b0 + b) * H + u0 + j;
    const float c = fg * c_prev[o] + ig * gg;
    c_out[o] = c;
    h_out[o] = og * tanhf(c);
    if (gates_out) {
        float* gp = gates_out + (int64_t)(b0 + b) * 4 * H + u0 + j;
        gp[0] = ig; gp[H] = fg; gp[2 * H] = gg; gp[3 * H] = og;
    }
}

// One BPTT step for UPB units x a 64-row batch tile.
//   dh = dhs_t + sum_j dgates_next[b,j] * W_hh[j,u]       (skipped at the last step)
//   do = dh*tanh(c); dc = dc_in + dh*o*(1-tanh(c)^2); di=dc*g; df=dc*c_prev; dg=dc*i
//   dgates_t = (di*i(1-i), df*f(1-f), dg*(1-g^2), do*o(1-o)); dc_out = dc*f
// 256 threads: 16-way split of each 64-wide K chunk x 16 row groups, 4x4 register tiles.
__global__ void __launch_bounds__(256)
lstm_step_bwd_kernel(const float* __restrict__ dhs_t, const float* __restrict__ dgates_next,
                     const float* __restrict__ w_hh, const float* __restrict__ gates_t,
                     const float* __restrict__ c_t, const float* __restrict__ c_prev,
                     float* __restrict__ dc_work, float* __restrict__ dgates_t, int first,
                     int B, int H) {
    __shared__ __align__(16) float Gs[LSTM_KC][LSTM_BT + 4];
    __shared__ __align__(16) float Ws[LSTM_KC][4];
    const int tid = threadIdx.x;
    const int u0 = blockIdx.x * LSTM_UPB, b0 = blockIdx.y * LSTM_BT;
    const int kq = tid >> 4, tb = tid & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int K = 4 * H;
    if (dgates_next) {
        for (int kc = 0; kc < K; kc += LSTM_KC) {
            const int k = tid & 63;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int b = (tid >> 6) + 4 * j;
                Gs[k][b] = (b0 + b < B && kc + k < K)
                               ? dgates_next[(int64_t)(b0 + b) * K + kc + k] : 0.f;
            }
            if (tid < LSTM_KC) {
                const float4 wv = (kc + tid < K)
                    ? __ldg(reinterpret_cast<const float4*>(w_hh + (int64_t)(kc + tid) * H + u0))
                    : make_float4(0.f, 0.f, 0.f, 0.f);
                *reinterpret_cast<float4*>(&Ws[tid][0]) = wv;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < LSTM_KC / 16; ++kk) {
                const int kx = kq * (LSTM_KC / 16) + kk;
                const float4 gv = *reinterpret_cast<const float4*>(&Gs[kx][tb * 4]);
                const float4 wv = *reinterpret_cast<const float4*>(&Ws[kx][0]);
                const float a[4] = {gv.x, gv.y, gv.z, gv.w}, w[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
    float* red = &Gs[0][0];  // [16][64][4] = 4096 floats <= 64*68
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) red[(kq * LSTM_BT + tb * 4 + i) * 4 + j] = acc[i][j];
    __syncthreads();
    const int b = tid >> 2, j = tid & 3;
    if (b0 + b >= B) return;
    float dh = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) dh += red[(q * LSTM_BT + b) * 4 + j];
    const int64_t o = (int64_t)(b0 + b) * H + u0 + j;
    dh += dhs_t[o];
    const float* gp = gates_t + (int64_t)(b0 + b) * 4 * H + u0 + j;
    const float ig = gp[0], fg = gp[H], gg = gp[2 * H], og = gp[3 * H];
    const float tc = tanhf(c_t[o]);
    const float dc_in = first ? 0.f : dc_work[o];
    const float dc = dc_in + dh * og * (1.f - tc * tc);
    float* dg = dgates_t + (int64_t)(b0 + b) * 4 * H + u0 + j;
    dg[0] = dc * gg * ig * (1.f - ig);
    dg[H] = dc * c_prev[o] * fg * (1.f - fg);
    dg[2 * H] = dc * ig * (1.f - gg * gg);
    dg[3 * H] = dh * tc * og * (1.f - og);
    dc_work[o] = dc * fg;
}

// ======================================================================== dueling head
// q[m,a] = val + adv_a - mean(adv); val = hid[m,:H].w2[0] + b2[0]; adv_a = hid[m,H:].w2[1+a] + b2[1+a]
__global__ void head_out_kernel(const float* __restrict__ hid, const float* __restrict__ w2,
                                const float* __restrict__ b2, float* __restrict__ q, int M,
                                int H) {
    const int lane = threadIdx.x & 31;
    const int m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (m >= M) return;
    const float* hv = hid + (int64_t)m * 2 * H;
    float s[7];
#pragma unroll
    for (int a = 0; a < 7; ++a) s[a] = 0.f;
    for (int j = lane; j < H; j += 32) {
        const float x = hv[j], y = hv[H + j];
        s[0] = fmaf(x, __ldg(w2 + j), s[0]);
#pragma unroll
        for (int a = 0; a < 6; ++a) s[1 + a] = fmaf(y, __ldg(w2 + (1 + a) * H + j), s[1 + a]);
    }
#pragma unroll
    for (int a = 0; a < 7; ++a) s[a] = warp_sum(s[a]);
    if (lane == 0) {
        const float val = s[0] + b2[0];
        float adv[6], mean = 0.f;
#pragma unroll
        for (int a = 0; a < 6; ++a) { adv[a] = s[1 + a] + b2[1 + a]; mean += adv[a]; }
        mean = mean / 6.f;
#pragma unroll
        for (int a = 0; a < 6; ++a) q[(int64_t)m * 6 + a] = val + adv[a] - mean;
    }
}

__device__ __forceinline__ void head_dz(const float* __restrict__ dq_row, float& dval,
                                        float dadv[6]) {
    dval = 0.f;
#pragma unroll
    for (int a = 0; a < 6; ++a) dval += dq_row[a];
    const float mean = dval / 6.f;
#pragma unroll
    for (int a = 0; a < 6; ++a) dadv[a] = dq_row[a] - mean;
}

// dhid[m,c] = relu'(hid) * (c<H ? dval*w2[0,c] : sum_a dadv_a*w2[1+a,c-H])
__global__ void head_bwd_dhid_kernel(const float* __restrict__ dq, const float* __restrict__ hid,
                                     const float* __restrict__ w2, float* __restrict__ dhid,
                                     int M, int H) {
    const int m = blockIdx.x;
    float dval, dadv[6];
    head_dz(dq + (int64_t)m * 6, dval, dadv);
    for (int c = threadIdx.x; c < 2 * H; c += blockDim.x) {
        float d;
        if (c < H) d = dval * __ldg(w2 + c);
        else {
            d = 0.f;
#pragma unroll
            for (int a = 0; a < 6; ++a) d = fmaf(dadv[a], __ldg(w2 + (1 + a) * H + c - H), d);
        }
        const int64_t o = (int64_t)m * 2 * H + c;
        dhid[o] = hid[o] > 0.f ? d : 0.f;
    }
}

// dw2[0,j] = sum_m dval_m hid[m,j]; dw2[1+a,j] = sum_m dadv_{m,a} hid[m,H+j]
// 32 columns per CTA, 32 warps stride over rows; block (H/32) also reduces db2.
__global__ void __launch_bounds__(1024)
head_bwd_dw2_kernel(const float* __restrict__ dq, const float* __restrict__ hid,
                    float* __restrict__ dw2, float* __restrict__ db2, int M, int H) {
    __shared__ float red[7][32][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const bool bias_block = (int)blockIdx.x == H / 32;
    const int j = blockIdx.x * 32 + lane;
    float s[7];
#pragma unroll
    for (int a = 0; a < 7; ++a) s[a] = 0.f;
    if (!bias_block) {
        for (int m = w; m < M; m += 32) {
            float dval, dadv[6];
            head_dz(dq + (int64_t)m * 6, dval, dadv);
            const float x = hid[(int64_t)m * 2 * H + j], y = hid[(int64_t)m * 2 * H + H + j];
            s[0] = fmaf(dval, x, s[0]);
#pragma unroll
            for (int a = 0; a < 6; ++a) s[1 + a] = fmaf(dadv[a], y, s[1 + a]);
        }
    } else {
        for (int m = threadIdx.x; m < M; m += 1024) {
            float dval, dadv[6];
            head_dz(dq + (int64_t)m * 6, dval, dadv);
            s[0] += dval;
#pragma unroll
            for (int a = 0; a < 6; ++a) s[1 + a] += dadv[a];
        }
    }
#pragma unroll
    for (int a = 0; a < 7; ++a) red[a][w][lane] = s[a];
    __syncthreads();
    if (!bias_block) {
        if (w < 7) {
            float t = 0.f;
            for (int i = 0; i < 32; ++i) t += red[w][i][lane];
            dw2[(int64_t)w * H + j] = t;
        }
    } else if (threadIdx.x < 7) {
        float t = 0.f;
        for (int i = 0; i < 32; ++i)
            for (int l = 0; l < 32; ++l) t += red[threadIdx.x][i][l];
        db2[threadIdx.x] = t;
    }
}

// ======================================================================== TD target (K7)
__device__ __forceinline__ float value_h(float x, float eps) {
    const float s = x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    return s * (sqrtf(fabsf(x) + 1.f) - 1.f) + eps * x;
}
__device__ __forceinline__ float value_h_inv(float x, float eps) {
    const float s = x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    const float t = (sqrtf(1.f + 4.f * eps * (fabsf(x) + 1.f + eps)) - 1.f) / (2.f * eps);
    return s * (t * t - 1.f);
}

// single CTA (n = L*B rows is a few thousand): deterministic loss reduction
__global__ void __launch_bounds__(1024)
td_target_kernel(const float* __restrict__ q_all, const float* __restrict__ qn_on,
                 const float* __restrict__ qn_tg, const int64_t* __restrict__ action,
                 const float* __restrict__ reward, const float* __restrict__ done, int n,
                 float gamma_n, int rescale, float rescale_eps, float prio_eps, float alpha,
                 const int64_t* __restrict__ na_override, int64_t* __restrict__ next_action,
                 float* __restrict__ y_out, float* __restrict__ delta_out,
                 float* __restrict__ dq, float* __restrict__ prio, float* __restrict__ loss) {
    __shared__ float red[32];
    float part = 0.f;
    const float inv_n = 1.f / (float)n;
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        const float* qo = qn_on + (int64_t)r * 6;
        int best = 0;
        float bv = qo[0];
#pragma unroll
        for (int a = 1; a < 6; ++a)
            if (qo[a] > bv) { bv = qo[a]; best = a; }   // first max wins, as torch.argmax
        if (na_override) best = (int)na_override[r];
        const float nq = qn_tg[(int64_t)r * 6 + best];
        const float nd = 1.f - done[r];
        float y;
        if (rescale) y = value_h(reward[r] + gamma_n * value_h_inv(nq, rescale_eps) * nd, rescale_eps);
        else y = __fadd_rn(reward[r], __fmul_rn(__fmul_rn(gamma_n, nq), nd));
        const int act = (int)action[r];
        const float d = q_all[(int64_t)r * 6 + act] - y;
        next_action[r] = best;
        y_out[r] = y;
        delta_out[r] = d;
        prio[r] = powf(fabsf(d) + prio_eps, alpha);
#pragma unroll
        for (int a = 0; a < 6; ++a) dq[(int64_t)r * 6 + a] = a == act ? d * inv_n : 0.f;
        part += 0.5f * d * d;
    }
    part = block_sum(part, red);
    if (threadIdx.x == 0) *loss = part * inv_n;
}

// ======================================================================== optimiser (K10)
__global__ void rmsprop_kernel(float* __restrict__ p, const float* __restrict__ g,
                               float* __restrict__ v, float* __restrict__ mavg, int64_t n,
                               float lr, float alpha, float oma, float eps, float gscale) {
    const int64_t n4 = n / 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (int64_t)gridDim.x * blockDim.x) {
        float4 P = reinterpret_cast<float4*>(p)[i];
        const float4 G0 = reinterpret_cast<const float4*>(g)[i];
        float4 V = reinterpret_cast<float4*>(v)[i];
        float4 Mv = reinterpret_cast<float4*>(mavg)[i];
        float pp[4] = {P.x, P.y, P.z, P.w}, gg[4] = {G0.x, G0.y, G0.z, G0.w};
        float vv[4] = {V.x, V.y, V.z, V.w}, mm[4] = {Mv.x, Mv.y, Mv.z, Mv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float gr = gg[j] * gscale;
            vv[j] = vv[j] * alpha + oma * gr * gr;
            mm[j] = mm[j] + oma * (gr - mm[j]);   // torch lerp_
            const float avg = sqrtf(vv[j] - mm[j] * mm[j]) + eps;
            pp[j] = pp[j] - lr * (gr / avg);
        }
        reinterpret_cast<float4*>(p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
        reinterpret_cast<float4*>(v)[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
        reinterpret_cast<float4*>(mavg)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    }
    if (blockIdx.x == 0) {
        for (int64_t i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
            const float gr = g[i] * gscale;
            const float vn = v[i] * alpha + oma * gr * gr;
            const float mn = mavg[i] + oma * (gr - mavg[i]);
            v[i] = vn; mavg[i] = mn;
            p[i] = p[i] - lr * (gr / (sqrtf(vn - mn * mn) + eps));
        }
    }
}

// ======================================================================== host-side pieces
template <class T>
static int torso_fwd_impl(const T* frames, FrameMap fm, int R, const float* w1, const float* b1,
                          const float* w2, const float* b2, const float* w3, const float* b3,
                          float* a1, float* a2, float* feat, cudaStream_t st) {
    using W = DenseLoad<true>;
    {
        ConvPatchRows<T, 4, 8, 84, 20, 4> al{frames, fm};
        W bl{w1, 256, 1};
        EpiConv<32, 400> ep{a1, b1};
        R2D2_CUDA_OK((launch_gemm<128, 32, 8, 4>("conv1_fwd", al, bl, ep, R * 400, 32, 256, 1, st)));
    }
    {
        ConvPatchRows<float, 32, 4, 20, 9, 2> al{a1, FrameMap{0, 0, 0, 12800}};
        W bl{w2, 512, 1};
        EpiConv<32, 81> ep{a2, b2};
        R2D2_CUDA_OK((launch_gemm<128, 32, 8, 4>("conv2_fwd", al, bl, ep, R * 81, 32, 512, 1, st)));
    }
    {
        ConvPatchRows<float, 32, 3, 9, 7, 1> al{a2, FrameMap{0, 0, 0, 2592}};
        W bl{w3, 288, 1};
        EpiConv<32, 49> ep{feat, b3};
        R2D2_CUDA_OK((launch_gemm<128, 32, 8, 4>("conv3_fwd", al, bl, ep, R * 49, 32, 288, 1, st)));
    }
    return R2D2_OK;
}

static int pick_splits(int tiles, int K) {
    int s = (2 * sm_count() * 4 + tiles - 1) / tiles;   // ~8 small CTAs per SM
    int maxs = K / 256;
    if (s > maxs) s = maxs;
    if (s < 1) s = 1;
    if (s > 512) s = 512;
    return s;
}

// conv weight gradient: dW[32, KP] = sum over (frame, out pixel)
template <class AL, class BL>
static int conv_wgrad(const char* name, const AL& al, const BL& bl, int KP, int Kred, float* dw, float* part,
                      cudaStream_t st) {
    const int tiles = (KP + 63) / 64;
    const int splits = pick_splits(tiles, Kred);
    EpiPartial ep{part, KP, (long long)32 * KP};
    R2D2_CUDA_OK((launch_gemm<32, 64, 4, 8>(name, al, bl, ep, 32, KP, Kred, splits, st)));
    const int64_t n = (int64_t)32 * KP;
    R2D2_LAUNCH("reduce_partials_kernel", st, reduce_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(part, dw, n, splits));
    return R2D2_OK;
}

constexpr int64_t WGRAD_PART_FLOATS = (int64_t)512 * 32 * 512;  // max splits x largest dW

template <class T>
static int torso_bwd_impl(const T* frames, FrameMap fm, int R, const float* a1, const float* a2,
                          const float* feat, float* dfeat, const float* w2, const float* w3,
                          float* dw1, float* db1, float* dw2, float* db2, float* dw3, float* db3,
                          void* scratch, cudaStream_t st, bool skip_wgrad = false) {
    R2D2_REQUIRE((int64_t)R * 400 < (1ll << 31), "torso_bwd: too many frames");
    float* dz2 = (float*)scratch;                   // [R,32,9,9]
    float* dz1 = dz2 + (int64_t)R * 2592;           // [R,32,20,20]
    float* part = dz1 + (int64_t)R * 12800;         // split-K partials
    {   // layer 3: dZ3 = dfeat * relu'(feat)
        const int64_t n = (int64_t)R * R2D2_FEAT;
        int64_t blocks = (n + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        R2D2_LAUNCH("relu_mask_kernel", st, relu_mask_kernel<<<(unsigned)blocks, 256, 0, st>>>(dfeat, feat, n));
    }
    {
        GradRowsT<32, 7> al{dfeat};
        ConvPatchCols<float, 32, 3, 9, 7, 1> bl{a2, FrameMap{0, 0, 0, 2592}};
        int rc = skip_wgrad ? 0 : conv_wgrad("conv3_wgrad", al, bl, 288, R * 49, dw3, part, st);
        if (rc) return rc;
        rc = chansum_dispatch(dfeat, db3, R, 49, part, st);
        if (rc) return rc;
    }
    {
        constexpr int smem = (32 * 3 * 3 * 32 + 32 * 49) * 4;
        R2D2_CUDA_OK(cudaFuncSetAttribute(conv_dgrad_kernel<3, 9, 7, 1>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        R2D2_LAUNCH("conv_dgrad_kernel<3,9,7,1>", st, conv_dgrad_kernel<3, 9, 7, 1><<<R, 256, smem, st>>>(dfeat, w3, a2, dz2));
    }
    {   // layer 2
        GradRowsT<32, 9> al{dz2};
        ConvPatchCols<float, 32, 4, 20, 9, 2> bl{a1, FrameMap{0, 0, 0, 12800}};
        int rc = skip_wgrad ? 0 : conv_wgrad("conv2_wgrad", al, bl, 512, R * 81, dw2, part, st);
        if (rc) return rc;
        rc = chansum_dispatch(dz2, db2, R, 81, part, st);
        if (rc) return rc;
    }
    {
        constexpr int smem = (32 * 4 * 4 * 32 + 32 * 81) * 4;
        R2D2_CUDA_OK(cudaFuncSetAttribute(conv_dgrad_kernel<4, 20, 9, 2>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        R2D2_LAUNCH("conv_dgrad_kernel<4,20,9,2>", st, conv_dgrad_kernel<4, 20, 9, 2><<<R, 256, smem, st>>>(dz2, w2, a1, dz1));
    }
    {   // layer 1 (no data gradient: frames are inputs)
        GradRowsT<32, 20> al{dz1};
        ConvPatchCols<T, 4, 8, 84, 20, 4> bl{frames, fm};
        int rc = skip_wgrad ? 0 : conv_wgrad("conv1_wgrad", al, bl, 256, R * 400, dw1, part, st);
        if (rc) return rc;
        rc = chansum_dispatch(dz1, db1, R, 400, part, st);
        if (rc) return rc;
    }
    return R2D2_OK;
}

}  // namespace r2d2

using namespace r2d2;

extern "C" {

int r2d2_torso_fwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* w1,
                      const float* b1, const float* w2, const float* b2, const float* w3,
                      const float* b3, float* a1, float* a2, float* feat, void* stream) {
    R2D2_REQUIRE(slab && w1 && b1 && w2 && b2 && w3 && b3 && a1 && a2 && feat, "torso_fwd: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && s0 >= 0 && s0 + nt <= S, "torso_fwd: bad frame window");
    R2D2_REQUIRE((int64_t)B * nt * 400 < (1ll << 31), "torso_fwd: too many frames");
    return torso_fwd_impl<uint8_t>(slab, FrameMap{B, S, s0, R2D2_FRAME_BYTES}, B * nt, w1, b1, w2,
                                   b2, w3, b3, a1, a2, feat, as_stream(stream));
}

int r2d2_torso_fwd_f32(const float* frames, int R, const float* w1, const float* b1,
                       const float* w2, const float* b2, const float* w3, const float* b3,
                       float* a1, float* a2, float* feat, void* stream) {
    R2D2_REQUIRE(frames && w1 && b1 && w2 && b2 && w3 && b3 && a1 && a2 && feat, "torso_fwd: null pointer");
    R2D2_REQUIRE(R > 0 && (int64_t)R * 400 < (1ll << 31), "torso_fwd: bad frame count");
    return torso_fwd_impl<float>(frames, FrameMap{0, 0, 0, R2D2_FRAME_BYTES}, R, w1, b1, w2, b2, w3,
                                 b3, a1, a2, feat, as_stream(stream));
}

int r2d2_linear_fwd(const float* x, const float* w, const float* bias1, const float* bias2,
                    float* y, int M, int N, int K, int relu, void* stream) {
    R2D2_REQUIRE(x && w && y && M > 0 && N > 0 && K > 0, "linear_fwd: bad args");
    DenseLoad<true> al{x, K, 1}, bl{w, K, 1};
    EpiDense ep{y, N, bias1, bias2, relu, 0};
    if (M >= 512) R2D2_CUDA_OK((launch_gemm<128, 128, 8, 8>("linear_fwd", al, bl, ep, M, N, K, 1, as_stream(stream))));
    else R2D2_CUDA_OK((launch_gemm<64, 64, 4, 4>("linear_fwd_small", al, bl, ep, M, N, K, 1, as_stream(stream))));
    return R2D2_OK;
}

int r2d2_linear_dgrad(const float* dy, const float* w, float* dx, int M, int N, int K,
                      int accumulate, void* stream) {
    R2D2_REQUIRE(dy && w && dx && M > 0 && N > 0 && K > 0, "linear_dgrad: bad args");
    DenseLoad<true> al{dy, N, 1};
    DenseLoad<false> bl{w, 1, K};
    EpiDense ep{dx, K, nullptr, nullptr, 0, accumulate};
    if (M >= 512) R2D2_CUDA_OK((launch_gemm<128, 128, 8, 8>("linear_dgrad", al, bl, ep, M, K, N, 1, as_stream(stream))));
    else R2D2_CUDA_OK((launch_gemm<64, 64, 4, 4>("linear_dgrad_small", al, bl, ep, M, K, N, 1, as_stream(stream))));
    return R2D2_OK;
}

static int wgrad_splits(int M, int N, int K) {
    const int tiles = ((N + 127) / 128) * ((K + 127) / 128);
    int s = (2 * 148 + tiles - 1) / tiles;
    int maxs = M / 128;
    if (s > maxs) s = maxs;
    if (s < 1) s = 1;
    return s;
}

int64_t r2d2_linear_wgrad_scratch_bytes(int M, int N, int K) {
    const int s = wgrad_splits(M, N, K);
    return s > 1 ? (int64_t)s * N * K * 4 : 0;
}

int r2d2_linear_wgrad(const float* dy, const float* x, float* dw, float* db, int M, int N,
                      int K, void* scratch, void* stream) {
    R2D2_REQUIRE(dy && x && dw && M > 0 && N > 0 && K > 0, "linear_wgrad: bad args");
    cudaStream_t st = as_stream(stream);
    DenseLoad<false> al{dy, 1, N}, bl{x, 1, K};
    const int splits = wgrad_splits(M, N, K);
    if (splits > 1) {
        R2D2_REQUIRE(scratch, "linear_wgrad: scratch required (%lld bytes)",
                     (long long)r2d2_linear_wgrad_scratch_bytes(M, N, K));
        EpiPartial ep{(float*)scratch, K, (long long)N * K};
        R2D2_CUDA_OK((launch_gemm<128, 128, 8, 8>("linear_wgrad", al, bl, ep, N, K, M, splits, st)));
        const int64_t n = (int64_t)N * K;
        R2D2_LAUNCH("reduce_partials_kernel", st, reduce_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const float*)scratch,
                                                                            dw, n, splits));
    } else {
        EpiDense ep{dw, K, nullptr, nullptr, 0, 0};
        R2D2_CUDA_OK((launch_gemm<128, 128, 8, 8>("linear_wgrad", al, bl, ep, N, K, M, 1, st)));
    }
    if (db) {
        R2D2_LAUNCH("colsum_kernel", st, colsum_kernel<<<(N + 31) / 32, 1024, 0, st>>>(dy, db, M, N));
    }
    return R2D2_OK;
}

int r2d2_lstm_seq_fwd(const float* xproj, const float* w_hh, float* hs_ext, float* cs_ext,
                      float* gates, int T, int B, int H, void* stream) {
    R2D2_REQUIRE(xproj && w_hh && hs_ext && cs_ext, "lstm_seq_fwd: null pointer");
    R2D2_REQUIRE(T > 0 && B > 0 && H > 0 && H % LSTM_UPB == 0, "lstm_seq_fwd: H must be a multiple of %d", LSTM_UPB);
    const int64_t bh = (int64_t)B * H;
    dim3 grid(H / LSTM_UPB, (B + LSTM_BT - 1) / LSTM_BT);
    for (int t = 0; t < T; ++t) {
        R2D2_LAUNCH("lstm_step_fwd_kernel", as_stream(stream), lstm_step_fwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(
            xproj + (int64_t)t * 4 * bh, w_hh, hs_ext + t * bh, cs_ext + t * bh,
            hs_ext + (t + 1) * bh, cs_ext + (t + 1) * bh, gates ? gates + (int64_t)t * 4 * bh : nullptr,
            B, H));
    }
    return R2D2_OK;
}

int r2d2_lstm_seq_bwd(const float* dhs, const float* w_hh, const float* gates,
                      const float* cs_ext, float* dgates, float* dc_work, int T, int B, int H,
                      void* stream) {
    R2D2_REQUIRE(dhs && w_hh && gates && cs_ext && dgates && dc_work, "lstm_seq_bwd: null pointer");
    R2D2_REQUIRE(T > 0 && B > 0 && H > 0 && H % LSTM_UPB == 0, "lstm_seq_bwd: bad sizes");
    R2D2_REQUIRE(((uintptr_t)w_hh % 16) == 0, "lstm_seq_bwd: w_hh must be 16-byte aligned");
    const int64_t bh = (int64_t)B * H;
    dim3 grid(H / LSTM_UPB, (B + LSTM_BT - 1) / LSTM_BT);
    for (int t = T - 1; t >= 0; --t) {
        R2D2_LAUNCH("lstm_step_bwd_kernel", as_stream(stream), lstm_step_bwd_kernel<<<grid, 256, 0, as_stream(stream)>>>(
            dhs + t * bh, t == T - 1 ? nullptr : dgates + (int64_t)(t + 1) * 4 * bh, w_hh,
            gates + (int64_t)t * 4 * bh, cs_ext + (t + 1) * bh, cs_ext + t * bh, dc_work,
            dgates + (int64_t)t * 4 * bh, t == T - 1 ? 1 : 0, B, H));
    }
    return R2D2_OK;
}

int r2d2_head_fwd(const float* hs, const float* w1, const float* b1, const float* w2,
                  const float* b2, float* hid, float* q, int M, int H, void* stream) {
    R2D2_REQUIRE(hs && w1 && b1 && w2 && b2 && hid && q && M > 0 && H > 0, "head_fwd: bad args");
    int rc = r2d2_linear_fwd(hs, w1, b1, nullptr, hid, M, 2 * H, H, 1, stream);
    if (rc) return rc;
    R2D2_LAUNCH("head_out_kernel", as_stream(stream), head_out_kernel<<<(M + 7) / 8, 256, 0, as_stream(stream)>>>(hid, w2, b2, q, M, H));
    return R2D2_OK;
}

int r2d2_reduce_partials(const float* part, float* out, int64_t n, int splits, void* stream) {
    R2D2_REQUIRE(part && out && n > 0 && splits > 0, "reduce_partials: bad args");
    R2D2_LAUNCH("reduce_partials_kernel", as_stream(stream),
                reduce_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, as_stream(stream)>>>(part, out, n, splits));
    return R2D2_OK;
}

int r2d2_head_out(const float* hid, const float* w2, const float* b2, float* q, int M, int H,
                  void* stream) {
    R2D2_REQUIRE(hid && w2 && b2 && q && M > 0 && H > 0, "head_out: bad args");
    R2D2_LAUNCH("head_out_kernel", as_stream(stream), head_out_kernel<<<(M + 7) / 8, 256, 0, as_stream(stream)>>>(hid, w2, b2, q, M, H));
    return R2D2_OK;
}

int r2d2_colsum(const float* dy, float* db, int M, int N, void* stream) {
    R2D2_REQUIRE(dy && db && M > 0 && N > 0, "colsum: bad args");
    R2D2_LAUNCH("colsum_kernel", as_stream(stream), colsum_kernel<<<(N + 31) / 32, 1024, 0, as_stream(stream)>>>(dy, db, M, N));
    return R2D2_OK;
}

int r2d2_colsum2(const float* dy, float* db, float* db_copy, int M, int N, void* stream) {
    R2D2_REQUIRE(dy && db && db_copy && M > 0 && N > 0, "colsum2: bad args");
    R2D2_LAUNCH("colsum_kernel", as_stream(stream), colsum_kernel<<<(N + 31) / 32, 1024, 0, as_stream(stream)>>>(dy, db, M, N, db_copy));
    return R2D2_OK;
}

int r2d2_head_bwd(const float* dq, const float* hid, const float* w2, float* dhid, float* dw2,
                  float* db2, int M, int H, void* stream) {
    R2D2_REQUIRE(dq && hid && w2 && M > 0 && H > 0 && H % 32 == 0,
                 "head_bwd: bad args (H must be a multiple of 32)");
    R2D2_REQUIRE(dhid || (dw2 && db2), "head_bwd: nothing to compute (dhid and dw2/db2 are NULL)");
    R2D2_REQUIRE((dw2 == nullptr) == (db2 == nullptr), "head_bwd: dw2 and db2 go together");
    // the two halves are independent: a caller may ask for one of them per call (dhid = NULL or
    // dw2 = db2 = NULL) to run them on different streams
    if (dhid)
        R2D2_LAUNCH("head_bwd_dhid_kernel", as_stream(stream), head_bwd_dhid_kernel<<<M, 256, 0, as_stream(stream)>>>(dq, hid, w2, dhid, M, H));
    if (dw2)
        R2D2_LAUNCH("head_bwd_dw2_kernel", as_stream(stream), head_bwd_dw2_kernel<<<H / 32 + 1, 1024, 0, as_stream(stream)>>>(dq, hid, dw2, db2, M, H));
    return R2D2_OK;
}

int r2d2_td_target(const float* q_all, const float* q_next_online, const float* q_next_target,
                   const int64_t* action, const float* reward, const float* done, int n,
                   float gamma_n, int rescale, float rescale_eps, float prio_eps, float alpha,
                   const int64_t* next_action_override, int64_t* next_action, float* y,
                   float* delta, float* dq, float* prio, float* loss, void* stream) {
    R2D2_REQUIRE(q_all && q_next_online && q_next_target && action && reward && done &&
                     next_action && y && delta && dq && prio && loss && n > 0,
                 "td_target: bad args");
    R2D2_LAUNCH("td_target_kernel", as_stream(stream), td_target_kernel<<<1, 1024, 0, as_stream(stream)>>>(
        q_all, q_next_online, q_next_target, action, reward, done, n, gamma_n, rescale, rescale_eps,
        prio_eps, alpha, next_action_override, next_action, y, delta, dq, prio, loss));
    return R2D2_OK;
}

// ------------------------------------------------------------------------------------------
// r2d2_linear_wgrad_static: r2d2_linear_wgrad for callers that have no way to hand us a
// workspace.  Some users of this C ABI are not PyTorch programs: an actor host or a serving
// process written in C/C++ that received every device buffer it owns from another component
// (sized for inputs and outputs only), that runs under a fixed HBM budget handed out entirely
// at start-up, or that sits on a latency-critical path where cudaMalloc is off limits because
// it synchronises the whole device and would stall the inference streams running next to it.
// For them the split-K partial sums live in the array below, which is part of the library
// image: every GPU that loads the library carries these 32 MiB until the process exits.
//
// Rules of the road
//   * Largest problem: wgrad_splits(M,N,K) * N * K floats must fit in
//     R2D2_STATIC_WGRAD_SCRATCH_BYTES; anything larger is refused with R2D2_EINVAL and a
//     message saying how much was needed (use r2d2_linear_wgrad with your own scratch then).
//   * The first call on each device looks up the array's device address (the name below is
//     only a host-side handle, never a pointer a kernel may dereference) with
//     cudaGetSymbolAddress while that device is current, and caches it.  Make that first
//     call during set-up, before any stream capture begins (a first call on a capturing
//     stream is refused).  `stream` must belong to the current device (checked).
//   * One array per device, shared by every call and every stream.  It still holds the
//     previous call's partial sums; that is harmless because each call overwrites every
//     element it later reads (all splits of all tiles are written, empty splits write
//     zeros), so nothing needs clearing.  Two calls whose kernels overlap on the GPU
//     (different streams, no ordering between them) race on it: serialise them.
//   * No cudaMalloc / cudaFree / allocating library call is made here, ever.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_static_wgrad_scratch[R2D2_STATIC_WGRAD_SCRATCH_BYTES / 4];

static int static_wgrad_scratch_ptr(float** out, cudaStream_t st) {
    static std::mutex mu;
    static float* ptrs[64] = {nullptr};
    int dev = -1;
    R2D2_CUDA_OK(cudaGetDevice(&dev));
    R2D2_REQUIRE(dev >= 0 && dev < 64, "linear_wgrad_static: device ordinal %d out of range", dev);
    // the array that gets used is the CURRENT device's copy: a stream of another device would run
    // the kernels there with this device's address
    int sdev = dev;
    R2D2_CUDA_OK(cudaStreamGetDevice(st, &sdev));
    R2D2_REQUIRE(sdev == dev, "linear_wgrad_static: the stream belongs to device %d but device %d is "
                 "current; make the stream's device current before the call", sdev, dev);
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        // first call on this device: the address look-up is not a capturable operation
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        R2D2_CUDA_OK(cudaStreamIsCapturing(st, &cap));
        R2D2_REQUIRE(cap == cudaStreamCaptureStatusNone,
                     "linear_wgrad_static: the first call on device %d looks up the built-in scratch "
                     "array and must be made before stream capture begins", dev);
        void* p = nullptr;
        R2D2_CUDA_OK(cudaGetSymbolAddress(&p, g_static_wgrad_scratch));
        R2D2_REQUIRE(p != nullptr, "linear_wgrad_static: cudaGetSymbolAddress returned null");
        ptrs[dev] = static_cast<float*>(p);
    }
    *out = ptrs[dev];
    return R2D2_OK;
}

int r2d2_linear_wgrad_static(const float* dy, const float* x, float* dw, float* db, int M, int N,
                             int K, void* stream) {
    R2D2_REQUIRE(dy && x && dw && M > 0 && N > 0 && K > 0, "linear_wgrad_static: bad args");
    cudaStream_t st = as_stream(stream);
    const int splits = wgrad_splits(M, N, K);
    const long long need = splits > 1 ? (long long)splits * N * K * 4 : 0;
    R2D2_REQUIRE(need <= (long long)R2D2_STATIC_WGRAD_SCRATCH_BYTES,
                 "linear_wgrad_static: M=%d N=%d K=%d needs %lld bytes of split-K scratch but the "
                 "built-in array holds %lld; call r2d2_linear_wgrad with caller-provided scratch",
                 M, N, K, need, (long long)R2D2_STATIC_WGRAD_SCRATCH_BYTES);
    using AL = DenseLoad<false>;
    AL al{dy, 1, N}, bl{x, 1, K};
    int kps = (M + splits - 1) / splits;
    kps = (kps + 15) / 16 * 16;
    const dim3 grid((N + 127) / 128, (K + 127) / 128, splits);
    if (splits > 1) {
        float* part = nullptr;
        int rc = static_wgrad_scratch_ptr(&part, st);
        if (rc) return rc;
        EpiPartial ep{part, K, (long long)N * K};
        R2D2_LAUNCH("linear_wgrad", st,
                    gemm_kernel<128, 128, 8, 8, AL, AL, EpiPartial><<<grid, 256, 0, st>>>(al, bl, ep, N, K, M, kps));
        const int64_t n = (int64_t)N * K;
        R2D2_LAUNCH("reduce_partials_kernel", st,
                    reduce_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(part, dw, n, splits));
    } else {
        EpiDense ep{dw, K, nullptr, nullptr, 0, 0};
        R2D2_LAUNCH("linear_wgrad", st,
                    gemm_kernel<128, 128, 8, 8, AL, AL, EpiDense><<<grid, 256, 0, st>>>(al, bl, ep, N, K, M, kps));
    }
    if (db) R2D2_LAUNCH("colsum_kernel", st, colsum_kernel<<<(N + 31) / 32, 1024, 0, st>>>(dy, db, M, N));
    return R2D2_OK;
}

int64_t r2d2_torso_bwd_scratch_bytes(int R) {
    return ((int64_t)R * (2592 + 12800) + WGRAD_PART_FLOATS) * 4;
}

int r2d2_torso_bwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                      const float* a2, const float* feat, float* dfeat, const float* w2,
                      const float* w3, float* dw1, float* db1, float* dw2, float* db2,
                      float* dw3, float* db3, void* scratch, void* stream) {
    R2D2_REQUIRE(slab && a1 && a2 && feat && dfeat && w2 && w3 && dw1 && db1 && dw2 && db2 && dw3 &&
                     db3 && scratch, "torso_bwd: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && s0 >= 0 && s0 + nt <= S, "torso_bwd: bad frame window");
    return torso_bwd_impl<uint8_t>(slab, FrameMap{B, S, s0, R2D2_FRAME_BYTES}, B * nt, a1, a2, feat,
                                   dfeat, w2, w3, dw1, db1, dw2, db2, dw3, db3, scratch,
                                   as_stream(stream));
}

// pieces of the conv backward the tensor-core path still takes from this file
int r2d2_relu_mask(float* d, const float* a, int64_t n, void* stream) {
    R2D2_REQUIRE(d && a && n > 0, "relu_mask: bad args");
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    R2D2_LAUNCH("relu_mask_kernel", as_stream(stream), relu_mask_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(d, a, n));
    return R2D2_OK;
}
// db[c] = sum over frames and pixels of dz[R, C=32, PIX]; part: R2D2_CHANSUM_PART_FLOATS of scratch
int r2d2_chansum(const float* dz, float* db, int R, int PIX, float* part, void* stream) {
    R2D2_REQUIRE(dz && db && part && R > 0 && PIX > 0, "chansum: bad args");
    return chansum_dispatch(dz, db, R, PIX, part, as_stream(stream));
}

// data gradients + bias gradients only (dz2/dz1 left in scratch for the tensor-core weight
// gradients, r2d2_tc_torso_wgrad_u8): scratch layout [dz2 | dz1 | partials] as r2d2_torso_bwd_u8
int r2d2_torso_bwd_dgrad_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                            const float* a2, const float* feat, float* dfeat, const float* w2,
                            const float* w3, float* db1, float* db2, float* db3, void* scratch,
                            void* stream) {
    R2D2_REQUIRE(slab && a1 && a2 && feat && dfeat && w2 && w3 && db1 && db2 && db3 && scratch,
                 "torso_bwd_dgrad: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && s0 >= 0 && s0 + nt <= S, "torso_bwd_dgrad: bad frame window");
    return torso_bwd_impl<uint8_t>(slab, FrameMap{B, S, s0, R2D2_FRAME_BYTES}, B * nt, a1, a2, feat,
                                   dfeat, w2, w3, nullptr, db1, nullptr, db2, nullptr, db3, scratch,
                                   as_stream(stream), true);
}

int r2d2_torso_bwd_f32(const float* frames, int R, const float* a1, const float* a2,
                       const float* feat, float* dfeat, const float* w2, const float* w3,
                       float* dw1, float* db1, float* dw2, float* db2, float* dw3, float* db3,
                       void* scratch, void* stream) {
    R2D2_REQUIRE(frames && a1 && a2 && feat && dfeat && w2 && w3 && dw1 && db1 && dw2 && db2 && dw3 &&
                     db3 && scratch, "torso_bwd: null pointer");
    R2D2_REQUIRE(R > 0, "torso_bwd: bad frame count");
    return torso_bwd_impl<float>(frames, FrameMap{0, 0, 0, R2D2_FRAME_BYTES}, R, a1, a2, feat, dfeat,
                                 w2, w3, dw1, db1, dw2, db2, dw3, db3, scratch, as_stream(stream));
}

int r2d2_rmsprop_centered(float* param, const float* grad, float* square_avg, float* grad_avg,
                          int64_t n, double lr, double alpha, double eps, double grad_scale,
                          void* stream) {
    R2D2_REQUIRE(param && grad && square_avg && grad_avg && n > 0, "rmsprop: bad args");
    R2D2_REQUIRE(((uintptr_t)param % 16) == 0 && ((uintptr_t)grad % 16) == 0 &&
                     ((uintptr_t)square_avg % 16) == 0 && ((uintptr_t)grad_avg % 16) == 0,
                 "rmsprop: buffers must be 16-byte aligned");
    int64_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks < 1) blocks = 1;
    R2D2_LAUNCH("rmsprop_kernel", as_stream(stream), rmsprop_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(param, grad, square_avg,
                                                                    grad_avg, n, (float)lr,
                                                                    (float)alpha, (float)(1.0 - alpha),
                                                                    (float)eps, (float)grad_scale));
    return R2D2_OK;
}

}  // extern "C"
