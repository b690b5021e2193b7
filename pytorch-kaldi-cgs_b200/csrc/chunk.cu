// Minibatch assembly from a chunk that is resident in HBM (the loop at core.py:183-200 of the reference:
// every sentence of the minibatch is copied into a zeroed [max_len, batch, dim] tensor behind a random number of leading
// zero rows).  The random offsets are host decisions (Python's `random`, drawn by the caller in the reference's order);
// the device does the byte movement: one pass that writes every output element once, zero padding included, so the
// separate torch.zeros fill and the per-sentence slice copies of the reference are one launch.
#include "common.cuh"

namespace hcgs {

// out[t][k][c] = chunk[beg[k] + t - left[k]][c] for left[k] <= t < left[k] + len[k], else 0.  One thread per four
// consecutive floats of the flat output when dim % 4 == 0 (16-byte accesses), else one float per thread.
template <int V>
__global__ void __launch_bounds__(256) assemble_batch_kernel(const float* __restrict__ chunk, int64_t chunk_rows, int dim,
                                                             const int64_t* __restrict__ beg,
                                                             const int32_t* __restrict__ len,
                                                             const int32_t* __restrict__ left, int batch,
                                                             int64_t n_vec, float* __restrict__ out) {
    const int dv = dim / V;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / dv;
        const int c = (int)(i - row * dv) * V;
        const int k = (int)(row % batch);
        const int64_t off = row / batch - (int64_t)__ldg(left + k);
        const int64_t src = __ldg(beg + k) + off;
        const bool in = off >= 0 && off < (int64_t)__ldg(len + k) && src >= 0 && src < chunk_rows;
        if (V == 4) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (in) v = __ldg(reinterpret_cast<const float4*>(chunk + src * dim + c));
            reinterpret_cast<float4*>(out)[i] = v;
        } else {
            out[i] = in ? __ldg(chunk + src * dim + c) : 0.f;
        }
    }
}

// The reverse movement at the other end of the path: rows (t, k) of a [t_max * batch, cols] result (row t * batch + k) of
// sentence k, t < len[k], are written to rows row_off[k] + t of a packed [sum(len), cols] matrix -- every sentence one
// contiguous [len, cols] block, padding rows dropped -- so that one device -> host copy moves exactly the matrices
// data_io.write_mat stores (core.py:238-249).
template <int V>
__global__ void __launch_bounds__(256) unpad_rows_kernel(const float* __restrict__ src, int64_t ld_src, int cols,
                                                         const int64_t* __restrict__ row_off,
                                                         const int32_t* __restrict__ len, int batch, int64_t n_vec,
                                                         float* __restrict__ dst) {
    const int cv = cols / V;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / cv;
        const int c = (int)(i - row * cv) * V;
        const int k = (int)(row % batch);
        const int64_t t = row / batch;
        if (t >= (int64_t)__ldg(len + k)) continue;
        const int64_t d = (__ldg(row_off + k) + t) * cols + c;
        if (V == 4)
            *reinterpret_cast<float4*>(dst + d) = __ldg(reinterpret_cast<const float4*>(src + row * ld_src + c));
        else
            dst[d] = __ldg(src + row * ld_src + c);
    }
}

}  // namespace hcgs

using namespace hcgs;

extern "C" int hcgs_assemble_batch(const float* chunk, int64_t chunk_rows, int dim, const int64_t* beg,
                                   const int32_t* len, const int32_t* left, int batch, int max_len, float* out,
                                   hcgs_stream stream) {
    HCGS_REQUIRE(chunk && beg && len && left && out, "hcgs_assemble_batch: NULL argument");
    HCGS_REQUIRE(chunk_rows >= 0 && dim >= 1 && batch >= 1 && max_len >= 0,
                 "hcgs_assemble_batch: bad shape (chunk_rows=%lld dim=%d batch=%d max_len=%d)", (long long)chunk_rows,
                 dim, batch, max_len);
    const int64_t total = (int64_t)max_len * batch * dim;
    if (total == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = dim % 4 == 0 && ((uintptr_t)chunk & 15) == 0 && ((uintptr_t)out & 15) == 0;
    const int64_t n_vec = vec ? total / 4 : total;
    int64_t blocks = (n_vec + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;   // eight resident CTAs of 256 threads per SM, grid-stride beyond
    if (vec)
        assemble_batch_kernel<4><<<(unsigned)blocks, 256, 0, st>>>(chunk, chunk_rows, dim, beg, len, left, batch, n_vec, out);
    else
        assemble_batch_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(chunk, chunk_rows, dim, beg, len, left, batch, n_vec, out);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_unpad_rows(const float* src, int64_t ld_src, int cols, const int64_t* row_off, const int32_t* len,
                               int batch, int t_max, float* dst, hcgs_stream stream) {
    HCGS_REQUIRE(src && row_off && len && dst, "hcgs_unpad_rows: NULL argument");
    HCGS_REQUIRE(cols >= 1 && ld_src >= cols && batch >= 1 && t_max >= 0, "hcgs_unpad_rows: bad shape");
    const int64_t total = (int64_t)t_max * batch * cols;
    if (total == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = cols % 4 == 0 && ld_src % 4 == 0 && ((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0;
    const int64_t n_vec = vec ? total / 4 : total;
    int64_t blocks = (n_vec + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (vec)
        unpad_rows_kernel<4><<<(unsigned)blocks, 256, 0, st>>>(src, ld_src, cols, row_off, len, batch, n_vec, dst);
    else
        unpad_rows_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(src, ld_src, cols, row_off, len, batch, n_vec, dst);
    HCGS_LAUNCH_CHECK();
    return 0;
}

// ---- guided HCGS on the device ------------------------------------------------------------------------
// guided_choices.guided_array_rows (guided_choices.py:4-30) ranks the column blocks of a block-row strip by the mean |W| of
// each block, computed with torch.nn.AvgPool2d on a CPU copy of the strip -- one device -> host copy and one pooling call per
// strip and level.  Here one launch pools every window of a level: a thread owns a window and adds its |w| in row-major
// order in fp32 and divides by the window area, which is the CPU kernel's own arithmetic (verified bit for bit,
// tests/test_guided_device.py), so the ranking -- still np.argsort on the few hundred means -- sees identical numbers.
namespace hcgs {

__global__ void __launch_bounds__(128) window_abs_mean_kernel(const float* __restrict__ w, int64_t ldw,
                                                              const int32_t* __restrict__ win, int n_win,
                                                              float* __restrict__ means) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_win) return;
    const int r0 = win[4 * i], r1 = win[4 * i + 1], c0 = win[4 * i + 2], c1 = win[4 * i + 3];
    float sum = 0.f;
    for (int r = r0; r < r1; ++r) {
        const float* row = w + (int64_t)r * ldw;
        for (int c = c0; c < c1; ++c) sum = __fadd_rn(sum, fabsf(__ldg(row + c)));
    }
    means[i] = __fdiv_rn(sum, (float)((int64_t)(r1 - r0) * (c1 - c0)));
}

__global__ void __launch_bounds__(128) fill_rects_kernel(float* __restrict__ out, int64_t ld,
                                                         const int32_t* __restrict__ rects, float value) {
    const int r0 = rects[4 * blockIdx.x], r1 = rects[4 * blockIdx.x + 1], c0 = rects[4 * blockIdx.x + 2],
              c1 = rects[4 * blockIdx.x + 3];
    const int wd = c1 - c0, n = (r1 - r0) * wd;
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[(int64_t)(r0 + i / wd) * ld + c0 + i % wd] = value;
}

}  // namespace hcgs

extern "C" int hcgs_window_abs_mean(const float* w, int64_t ldw, const int32_t* windows, int n_windows, float* means,
                                    hcgs_stream stream) {
    HCGS_REQUIRE(w && windows && means, "hcgs_window_abs_mean: NULL argument");
    if (n_windows <= 0) return 0;
    hcgs::window_abs_mean_kernel<<<(n_windows + 127) / 128, 128, 0, (cudaStream_t)stream>>>(w, ldw, windows, n_windows, means);
    HCGS_LAUNCH_CHECK();
    return 0;
}

extern "C" int hcgs_fill_rects(float* out, int64_t ld, const int32_t* rects, int n_rects, float value, hcgs_stream stream) {
    HCGS_REQUIRE(out && rects, "hcgs_fill_rects: NULL argument");
    if (n_rects <= 0) return 0;
    hcgs::fill_rects_kernel<<<n_rects, 128, 0, (cudaStream_t)stream>>>(out, ld, rects, value);
    HCGS_LAUNCH_CHECK();
    return 0;
}
