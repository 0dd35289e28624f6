// Motion-aware terrain post-processing: which heightfield cells a motion touches.
// terrain_util.compute_hf_mask_inds (PARC/util/terrain_util.py:1851-1897) is a Python loop over frames x bodies x
// points; here one block per frame marks the cells of its <= 2048 body sample points in a shared-memory bitmap of
// the grid, and reads the bitmap back in index order -- which IS the sorted unique list torch.unique(dim=0) returns
// (key = ix * Y + iy is lexicographic in (ix, iy)). The lowest point height per cell is an ordered-int atomic min.
#include "parc_common.cuh"
#include "parc_lanes.cuh"
#include "../../include/parc_b200_terrain.h"

namespace parc {

// float min through integer atomics (values of either sign)
__device__ __forceinline__ void atomic_min_float(float* addr, float v) {
    v += 0.0f;                                           // -0.0 -> +0.0: its bit pattern (INT_MIN) would win every signed min
    if (v >= 0.f) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMax(reinterpret_cast<unsigned*>(addr), __float_as_uint(v));
}

__global__ void fill_kernel(float* __restrict__ p, long long n, float v, unsigned char* __restrict__ m) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
        p[k] = v;
        if (m) m[k] = 0;
    }
}

constexpr int MASK_THREADS = 256;
__global__ void __launch_bounds__(MASK_THREADS) hf_mask_inds_kernel(
    const int J, const __grid_constant__ parc_terrain_t tr, const float* __restrict__ body_pos, const float* __restrict__ body_rot,
    const float* __restrict__ body_points, const int* __restrict__ point_start, const int Np, long long* __restrict__ inds,
    int* __restrict__ counts, float* __restrict__ min_heights, unsigned char* __restrict__ hf_mask) {
    extern __shared__ unsigned s_bits[];             // ceil(X*Y / 32) words
    __shared__ int s_pstart[PARC_MAX_BODIES + 1];
    __shared__ int s_warp[MASK_THREADS / 32];
    __shared__ int s_base;
    const long long t = blockIdx.x;
    const int tid = threadIdx.x;
    const int cells = tr.dim_x * tr.dim_y, words = (cells + 31) >> 5;
    for (int w = tid; w < words; w += MASK_THREADS) s_bits[w] = 0u;
    for (int b = tid; b <= J; b += MASK_THREADS) s_pstart[b] = point_start[b];
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int k = tid; k < Np; k += MASK_THREADS) {
        int body = 0;
        while (body + 1 < J && s_pstart[body + 1] <= k) ++body;
        const float* bp = body_pos + (t * J + body) * 3;
        const Quat rot = quat_from_f4(__ldg(reinterpret_cast<const float4*>(body_rot) + t * J + body));
        const Vec3 lp = make_vec3(__ldg(body_points + k * 3), __ldg(body_points + k * 3 + 1), __ldg(body_points + k * 3 + 2));
        const Vec3 P = quat_rotate(rot, lp) + make_vec3(__ldg(bp), __ldg(bp + 1), __ldg(bp + 2));          // :1880
        const int ix = grid_index(P.x, tr.min_x, tr.dx, tr.dim_x), iy = grid_index(P.y, tr.min_y, tr.dy, tr.dim_y);   // :1882
        const int key = ix * tr.dim_y + iy;
        atomicOr(&s_bits[key >> 5], 1u << (key & 31));
        atomic_min_float(min_heights + key, P.z);                                                           // :1893-1896
        if (hf_mask) hf_mask[key] = 1;
    }
    __syncthreads();
    // ordered read-back: chunks of MASK_THREADS words, exclusive scan of the popcounts inside a chunk, running base
    long long* out = inds + t * (long long)Np * 2;
    for (int w0 = 0; w0 < words; w0 += MASK_THREADS) {
        const int w = w0 + tid;
        const unsigned bits = w < words ? s_bits[w] : 0u;
        const int cnt = __popc(bits);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL_MASK, incl, o); if ((tid & 31) >= o) incl += v; }
        if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
        __syncthreads();
        int warp_off = 0, total = 0;
        for (int k = 0; k < MASK_THREADS / 32; ++k) { const int v = s_warp[k]; if (k < (tid >> 5)) warp_off += v; total += v; }
        int pos = s_base + warp_off + incl - cnt;
        unsigned b = bits;
        while (b) {
            const int bit = __ffs(b) - 1;
            b &= b - 1;
            const int key = (w << 5) + bit;
            out[pos * 2] = key / tr.dim_y; out[pos * 2 + 1] = key % tr.dim_y;
            ++pos;
        }
        __syncthreads();
        if (tid == 0) s_base += total;
        __syncthreads();
    }
    if (tid == 0) counts[t] = s_base;
}

__global__ void hf_extra_vals_kernel(const __grid_constant__ parc_terrain_t tr, const unsigned char* __restrict__ mask,
                                     const float* __restrict__ min_heights, float hi, float lo, float jump_buf, float* __restrict__ mm) {
    const long long cells = (long long)tr.dim_x * tr.dim_y;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < cells; k += (long long)gridDim.x * blockDim.x) {
        const float h = __ldg(tr.hf + k);
        float a = hi, b = lo;                                             // :1932-1933
        if (mask[k]) {
            a = h; b = h;                                                 // :1935-1937
            const float mb = __ldg(min_heights + k);
            if (__fsub_rn(mb, h) >= jump_buf) { a = __fsub_rn(mb, jump_buf); b = lo; }   // :1940-1945
        }
        mm[k * 2] = a; mm[k * 2 + 1] = b;
    }
}

}  // namespace parc

using namespace parc;

extern "C" int parc_b200_hf_mask_inds(const parc_char_model_t* m, const parc_terrain_t* tr, const float* body_pos, const float* body_rot,
                                      int64_t T, const float* body_points, const int32_t* point_start, int32_t num_points,
                                      int64_t* inds, int32_t* counts, float* min_body_heights, uint8_t* hf_mask, void* stream) {
    if (!model_ok(m, "hf_mask_inds")) return PARC_ERR_INVALID_ARG;
    PARC_REQUIRE(tr && tr->hf && tr->dim_x > 0 && tr->dim_y > 0, "hf_mask_inds: empty terrain");
    PARC_REQUIRE(T >= 0, "hf_mask_inds: negative frame count");
    PARC_REQUIRE(min_body_heights != nullptr, "hf_mask_inds: min_body_heights is required");
    const long long cells = (long long)tr->dim_x * tr->dim_y;
    PARC_REQUIRE(cells <= 1600000, "hf_mask_inds: %lld cells exceed the shared-memory bitmap (1.6 M)", cells);
    fill_kernel<<<(unsigned)std::min<long long>((cells + 255) / 256, kNumSMs * 8), 256, 0, as_stream(stream)>>>(min_body_heights, cells,
                                                                                                               99999.9999f, hf_mask);
    int rc = check_launch("hf_mask_inds (fill)");
    if (rc || T == 0) return rc;
    PARC_REQUIRE(body_pos && body_rot && body_points && point_start && inds && counts, "hf_mask_inds: NULL pointer");
    PARC_REQUIRE(num_points > 0 && num_points <= 2048, "hf_mask_inds: %d body sample points (1..2048 supported)", num_points);
    const size_t smem = sizeof(unsigned) * (size_t)((cells + 31) / 32);
    static size_t smem_attr_of_device[64] = {};            // the opt-in is a per-device function attribute
    int dev = 0;
    cudaGetDevice(&dev);
    size_t& smem_attr = smem_attr_of_device[dev & 63];
    if (smem_attr == 0) smem_attr = 48 * 1024;
    if (smem > smem_attr) {
        cudaError_t ce = cudaFuncSetAttribute(hf_mask_inds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ce != cudaSuccess) { set_error("hf_mask_inds: %s", cudaGetErrorString(ce)); return PARC_ERR_CUDA; }
        smem_attr = smem;
    }
    hf_mask_inds_kernel<<<(unsigned)T, MASK_THREADS, smem, as_stream(stream)>>>(m->num_bodies, *tr, body_pos, body_rot, body_points, point_start,
                                                                             num_points, (long long*)inds, counts, min_body_heights, hf_mask);
    return check_launch("hf_mask_inds");
}

extern "C" int parc_b200_hf_extra_vals(const parc_terrain_t* tr, const uint8_t* hf_mask, const float* min_body_heights, float max_h,
                                       float min_h, float z_buf, float jump_buf, float* hf_maxmin, void* stream) {
    PARC_REQUIRE(tr && tr->hf && tr->dim_x > 0 && tr->dim_y > 0, "hf_extra_vals: empty terrain");
    PARC_REQUIRE(hf_mask && min_body_heights && hf_maxmin, "hf_extra_vals: NULL pointer");
    const long long cells = (long long)tr->dim_x * tr->dim_y;
    // max_h + z_buf and min_h - z_buf as the reference's python-float arithmetic rounds them when stored to fp32 (:1932-1933)
    const float hi = (float)((double)max_h + (double)z_buf), lo = (float)((double)min_h - (double)z_buf);
    hf_extra_vals_kernel<<<(unsigned)std::min<long long>((cells + 255) / 256, kNumSMs * 8), 256, 0, as_stream(stream)>>>(
        *tr, hf_mask, min_body_heights, hi, lo, jump_buf, hf_maxmin);
    return check_launch("hf_extra_vals");
}
