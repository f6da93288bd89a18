// nb_aux.cu -- O(N) helper kernels around the force kernels:
//   pack / unpack between the host packing of the reference
//   (pos_xyzm[4i..], vel_xyz[3i..]; cuda/exhaustive.cu:240-251) and the device
//   layouts (nb_direct.cuh), the stand-alone kick+drift used when a(x_k) is
//   already stored (serial/exhaustive.c:221-232), and get_energy
//   (common/get_energy.c:7-47) as an O(N^2) FP64 kernel.
#include "nb_aux.cuh"
#include <mutex>
#include "../../include/nbody_engine.h"

namespace nb {
namespace {

// Are all masses equal (as floats for an FP32 engine, as doubles for an FP64 one) and positive?
// cls[0] starts at 1 (mass_class below) and is cleared by any thread that sees a different
// mass; cls[1] = bits of (float)m_0, cls[2..3] = bits of m_0.  The direct kernels then drop m_j
// from their inner loops (nb_direct_f32.cu MODE 2, nb_direct_f64.cu UNI).  Runs on the staged
// host packing, so the host never walks the masses itself.
template <bool F64>
__global__ void mass_class_kernel(const double *__restrict__ pos, unsigned n, unsigned *__restrict__ cls)
{
    const double d0 = pos[3];
    const float m0 = (float)d0;
    bool differ = false;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double m = pos[4 * (size_t)i + 3];
        differ |= F64 ? (m != d0) : ((float)m != m0);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        cls[1] = __float_as_uint(m0);
        cls[2] = (unsigned)__double2loint(d0);
        cls[3] = (unsigned)__double2hiint(d0);
        // finite and positive; FP64: G*m and m / eps2^1.5 must stay ordinary numbers
        differ |= F64 ? !(d0 > 1e-200 && d0 < 1e200) : !(m0 > 0.f);
    }
    if (__any_sync(0xffffffffu, differ) && (threadIdx.x & 31) == 0) cls[0] = 0u;
}

__global__ void pack_f32_kernel(const double *__restrict__ pos, const double *__restrict__ vel,
                                unsigned n, unsigned n_quads, float4 *__restrict__ q0,
                                float4 *__restrict__ q1, float4 *__restrict__ v,
                                const unsigned *__restrict__ cls)
{
    const unsigned iq = blockIdx.x * blockDim.x + threadIdx.x;
    if (iq >= n_quads) return;
    // equal-mass kernels ignore m_j: their padding is parked where r^-3 underflows to exactly 0
    const float pad_pos = (cls && cls[0]) ? 1.0e18f : 0.f;
    float c[7][4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const unsigned b = iq * 4 + r;
        const bool real = b < n;
        // padding: massless at the origin, or (uniform-mass kernels, which ignore m_j) so far
        // away that r^-3 underflows to exactly 0
        c[0][r] = real ? (float)pos[4 * (size_t)b + 0] : pad_pos;
        c[1][r] = real ? (float)pos[4 * (size_t)b + 1] : pad_pos;
        c[2][r] = real ? (float)pos[4 * (size_t)b + 2] : pad_pos;
        c[3][r] = real ? (float)pos[4 * (size_t)b + 3] : 0.f;
        c[4][r] = (real && vel) ? (float)vel[3 * (size_t)b + 0] : 0.f;
        c[5][r] = (real && vel) ? (float)vel[3 * (size_t)b + 1] : 0.f;
        c[6][r] = (real && vel) ? (float)vel[3 * (size_t)b + 2] : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float4 f = make_float4(c[k][0], c[k][1], c[k][2], c[k][3]);
        q0[(size_t)iq * 4 + k] = f;
        q1[(size_t)iq * 4 + k] = f;
    }
    if (v) {
#pragma unroll
        for (int k = 0; k < 3; ++k)
            v[(size_t)iq * 3 + k] = make_float4(c[4 + k][0], c[4 + k][1], c[4 + k][2], c[4 + k][3]);
    }
}

__global__ void pack_f64_kernel(const double *__restrict__ pos, const double *__restrict__ vel,
                                unsigned n, unsigned n_pad, double4 *__restrict__ q0,
                                double4 *__restrict__ q1, double4 *__restrict__ v,
                                const unsigned *__restrict__ cls)
{
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_pad) return;
    const bool real = b < n;
    // padding: massless at the origin, or (equal-mass kernel, which ignores m_j) at 1e150, where
    // r2 = 3e300 is finite and y0^3 = 1.9e-451 underflows to exactly 0
    const double pad = (cls && cls[0]) ? 1.0e150 : 0.0;
    const double4 p = real ? make_double4(pos[4 * (size_t)b], pos[4 * (size_t)b + 1],
                                          pos[4 * (size_t)b + 2], pos[4 * (size_t)b + 3])
                           : make_double4(pad, pad, pad, 0);
    q0[b] = p;
    q1[b] = p;
    if (v)
        v[b] = (real && vel) ? make_double4(vel[3 * (size_t)b], vel[3 * (size_t)b + 1],
                                            vel[3 * (size_t)b + 2], 0.0)
                             : make_double4(0, 0, 0, 0);
}

// planes = 4: positions+mass -> out[4b..]; planes = 3: vel / acc -> out[3b..]
__global__ void unpack_f32_kernel(const float4 *__restrict__ src, unsigned n, int planes,
                                  double *__restrict__ out, const unsigned *__restrict__ scatter)
{
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    const float *f = reinterpret_cast<const float *>(src) + (size_t)(b >> 2) * (planes * 4) + (b & 3);
    const size_t row = scatter ? scatter[b] : b;
    for (int k = 0; k < planes; ++k) out[(size_t)planes * row + k] = (double)f[k * 4];
}

__global__ void unpack_f64_kernel(const double4 *__restrict__ src, unsigned n, int planes,
                                  double *__restrict__ out, const unsigned *__restrict__ scatter)
{
    const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    const double4 p = src[b];
    const size_t row = scatter ? scatter[b] : b;
    out[(size_t)planes * row + 0] = p.x;
    out[(size_t)planes * row + 1] = p.y;
    out[(size_t)planes * row + 2] = p.z;
    if (planes == 4) out[(size_t)planes * row + 3] = p.w;
}

__global__ void remap_chunks_kernel(const unsigned *__restrict__ in, unsigned *__restrict__ out, unsigned n,
                                    unsigned chunk, const unsigned *__restrict__ chunk_base)
{
    const unsigned m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < n) out[chunk_base[m / chunk] + m % chunk] = in[m];
}

__global__ void permute_rows_kernel(const double *__restrict__ src, double *__restrict__ dst,
                                    const unsigned *__restrict__ order, unsigned n, int planes)
{
    const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const size_t from = order[k];
    for (int c = 0; c < planes; ++c) dst[(size_t)planes * k + c] = src[(size_t)planes * from + c];
}

__global__ void kickdrift_f32_kernel(const float4 *__restrict__ qc, float4 *__restrict__ qn,
                                     float4 *__restrict__ vel, const float4 *__restrict__ acc,
                                     unsigned i_begin, unsigned i_end, unsigned n_real, float dt)
{
    const unsigned iq = i_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (iq >= i_end) return;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float4 X = qc[(size_t)iq * 4 + c], V = vel[(size_t)iq * 3 + c],
                     A = acc[(size_t)iq * 3 + c];
        float x[4] = {X.x, X.y, X.z, X.w}, v[4] = {V.x, V.y, V.z, V.w};
        const float a[4] = {A.x, A.y, A.z, A.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            if (iq * 4u + r >= n_real) continue;
            v[r] = __fadd_rn(v[r], __fmul_rn(__fmul_rn(a[r], dt), 0.5f));
            x[r] = __fadd_rn(x[r], __fmul_rn(v[r], dt));
        }
        vel[(size_t)iq * 3 + c] = make_float4(v[0], v[1], v[2], v[3]);
        qn[(size_t)iq * 4 + c] = make_float4(x[0], x[1], x[2], x[3]);
    }
}

__global__ void kickdrift_f64_kernel(const double4 *__restrict__ qc, double4 *__restrict__ qn,
                                     double4 *__restrict__ vel, const double4 *__restrict__ acc,
                                     unsigned i_begin, unsigned i_end, unsigned n_real, double dt)
{
    const unsigned b = i_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= i_end || b >= n_real) return;
    double4 X = qc[b], V = vel[b];
    const double4 A = acc[b];
    V.x = __dadd_rn(V.x, __dmul_rn(__dmul_rn(A.x, dt), 0.5));
    V.y = __dadd_rn(V.y, __dmul_rn(__dmul_rn(A.y, dt), 0.5));
    V.z = __dadd_rn(V.z, __dmul_rn(__dmul_rn(A.z, dt), 0.5));
    X.x = __dadd_rn(X.x, __dmul_rn(V.x, dt));
    X.y = __dadd_rn(X.y, __dmul_rn(V.y, dt));
    X.z = __dadd_rn(X.z, __dmul_rn(V.z, dt));
    vel[b] = V;
    qn[b] = X;
}

// get_energy: KE = 1/2 sum m v^2 ; PE = -G sum_{i<j} m_i m_j / r_ij (r = 0 pairs
// contribute 0, common/get_energy.c:38-40).  One thread per i, j > i tiles in
// shared memory, block partials written out and added on the host in order.
constexpr int ETB = 256;

// 1/r from MUFU.RSQ64H + one cubically convergent step (relative error < 2^-58); 0 for r = 0
__device__ __forceinline__ double inv_r(double r2)
{
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(r2));
    const double e = fma(-r2 * y0, y0, 1.0);
    const double y = fma(y0 * e, fma(0.375, e, 0.5), y0);
    return r2 > 0.0 ? y : 0.0;
}

// Rows of tile t against all j > i.  Row tiles near the top of the triangle have many
// columns, those near the bottom few: a CTA takes tile b and tile (last - b), so all CTAs
// carry the same number of pairs.
__global__ void __launch_bounds__(ETB)
    energy_kernel(const double *__restrict__ pos, const double *__restrict__ vel, unsigned n,
                  double *__restrict__ partial /* [grid][2] */)
{
    __shared__ double4 tile[ETB];
    __shared__ double red[2][ETB / 32];
    const unsigned ntiles = (n + ETB - 1) / ETB;
    double ke = 0, pe = 0;
    for (int half = 0; half < 2; ++half) {
        const unsigned t = half == 0 ? blockIdx.x : ntiles - 1 - blockIdx.x;
        if (half == 1 && t <= blockIdx.x) break; // odd tile count: the middle tile is done once
        const unsigned i = t * ETB + threadIdx.x;
        const bool real = i < n;
        double xi = 0, yi = 0, zi = 0, mi = 0, pei = 0;
        if (real) {
            xi = pos[4 * (size_t)i]; yi = pos[4 * (size_t)i + 1]; zi = pos[4 * (size_t)i + 2];
            mi = pos[4 * (size_t)i + 3];
            const double vx = vel[3 * (size_t)i], vy = vel[3 * (size_t)i + 1], vz = vel[3 * (size_t)i + 2];
            ke += mi * (vx * vx + vy * vy + vz * vz);
        }
        for (unsigned j0 = t * ETB; j0 < n; j0 += ETB) {
            const unsigned j = j0 + threadIdx.x;
            __syncthreads();
            tile[threadIdx.x] = j < n ? make_double4(pos[4 * (size_t)j], pos[4 * (size_t)j + 1],
                                                     pos[4 * (size_t)j + 2], pos[4 * (size_t)j + 3])
                                      : make_double4(0, 0, 0, 0);
            __syncthreads();
            const unsigned lim = min((unsigned)ETB, n - j0);
            if (j0 == t * ETB) { // diagonal tile: j > i only
                for (unsigned k = threadIdx.x + 1; k < lim; ++k) {
                    const double4 b = tile[k];
                    const double dx = b.x - xi, dy = b.y - yi, dz = b.z - zi;
                    pei = fma(b.w, inv_r(fma(dz, dz, fma(dy, dy, dx * dx))), pei);
                }
            } else {
#pragma unroll 4
                for (unsigned k = 0; k < lim; ++k) {
                    const double4 b = tile[k];
                    const double dx = b.x - xi, dy = b.y - yi, dz = b.z - zi;
                    pei = fma(b.w, inv_r(fma(dz, dz, fma(dy, dy, dx * dx))), pei);
                }
            }
        }
        if (real) pe += mi * pei;
    }
    for (int o = 16; o > 0; o >>= 1) {
        ke += __shfl_xor_sync(0xffffffffu, ke, o);
        pe += __shfl_xor_sync(0xffffffffu, pe, o);
    }
    if ((threadIdx.x & 31) == 0) {
        red[0][threadIdx.x >> 5] = ke;
        red[1][threadIdx.x >> 5] = pe;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int w = 0; w < ETB / 32; ++w) { a += red[0][w]; b += red[1][w]; }
        partial[2 * blockIdx.x] = a;
        partial[2 * blockIdx.x + 1] = b;
    }
}

inline unsigned cdiv(unsigned a, unsigned b) { return (a + b - 1) / b; }
} // namespace

cudaError_t mass_class(const double *pos, unsigned n, unsigned *cls, bool enabled, bool f64,
                       cudaStream_t s)
{
    const unsigned init[4] = {enabled && n ? 1u : 0u, 0u, 0u, 0u};
    cudaError_t e = cudaMemcpyAsync(cls, init, sizeof(init), cudaMemcpyHostToDevice, s); // pageable: staged at once
    if (e != cudaSuccess) return e;
    if (enabled && n) {
        const unsigned g = cdiv(n, 256) < 1184u ? cdiv(n, 256) : 1184u;
        if (f64) mass_class_kernel<true><<<g, 256, 0, s>>>(pos, n, cls);
        else mass_class_kernel<false><<<g, 256, 0, s>>>(pos, n, cls);
    }
    return cudaGetLastError();
}

cudaError_t pack_bodies(int precision, const double *pos, const double *vel, unsigned n,
                        unsigned n_units, void *q0, void *q1, void *v, const unsigned *cls, cudaStream_t s)
{
    if (precision == 1)
        pack_f32_kernel<<<cdiv(n_units, 256), 256, 0, s>>>(pos, vel, n, n_units, (float4 *)q0,
                                                          (float4 *)q1, (float4 *)v, cls);
    else
        pack_f64_kernel<<<cdiv(n_units, 256), 256, 0, s>>>(pos, vel, n, n_units, (double4 *)q0,
                                                          (double4 *)q1, (double4 *)v, cls);
    return cudaGetLastError();
}

cudaError_t unpack_bodies(int precision, const void *src, unsigned n, int planes, double *out,
                          const unsigned *scatter, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    if (precision == 1)
        unpack_f32_kernel<<<cdiv(n, 256), 256, 0, s>>>((const float4 *)src, n, planes, out, scatter);
    else
        unpack_f64_kernel<<<cdiv(n, 256), 256, 0, s>>>((const double4 *)src, n, planes, out, scatter);
    return cudaGetLastError();
}

cudaError_t remap_chunks(const unsigned *in, unsigned *out, unsigned n, unsigned chunk,
                         const unsigned *chunk_base, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    remap_chunks_kernel<<<cdiv(n, 256), 256, 0, s>>>(in, out, n, chunk, chunk_base);
    return cudaGetLastError();
}

cudaError_t permute_rows(const double *src, double *dst, const unsigned *order, unsigned n,
                         int planes, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    permute_rows_kernel<<<cdiv(n, 256), 256, 0, s>>>(src, dst, order, n, planes);
    return cudaGetLastError();
}

cudaError_t kickdrift(int precision, const void *qc, void *qn, void *vel, const void *acc,
                      unsigned i_begin, unsigned i_end, unsigned n_real, double dt, cudaStream_t s)
{
    if (i_end <= i_begin) return cudaSuccess;
    const unsigned cnt = i_end - i_begin;
    if (precision == 1)
        kickdrift_f32_kernel<<<cdiv(cnt, 256), 256, 0, s>>>((const float4 *)qc, (float4 *)qn,
                                                           (float4 *)vel, (const float4 *)acc,
                                                           i_begin, i_end, n_real, (float)dt);
    else
        kickdrift_f64_kernel<<<cdiv(cnt, 256), 256, 0, s>>>((const double4 *)qc, (double4 *)qn,
                                                           (double4 *)vel, (const double4 *)acc,
                                                           i_begin, i_end, n_real, dt);
    return cudaGetLastError();
}

unsigned energy_blocks(unsigned n) { return cdiv(cdiv(n, ETB), 2); } // a CTA takes two row tiles

cudaError_t energy_partials(const double *pos_xyzm, const double *vel_xyz, unsigned n,
                            double *partial, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    energy_kernel<<<energy_blocks(n), ETB, 0, s>>>(pos_xyzm, vel_xyz, n, partial);
    return cudaGetLastError();
}

// ------------------------------------------------ energy without allocations ----
// nb_energy_static (include/nbody_engine.h): energy_kernel above with its block partials in a
// file-scope device array instead of the engine's cudaMalloc'ed buffer, for callers that may not
// allocate (stream capture, serving loops).  Every GPU that loads the library holds the array
// (32 KB + 256 B) until exit.  All calls on a device share it and it keeps the last call's
// contents: a call overwrites the partials of every block it launches and reads only those, so
// nothing has to be cleared -- but two calls that overlap on the GPU (different streams) race.
constexpr unsigned kEnergyStaticMaxBlocks = NB_ENERGY_STATIC_MAX_BODIES / ETB / 2;
static __device__ __align__(256) double g_energy_static_partial[2 * kEnergyStaticMaxBlocks];

namespace {
// partials added in block order, like nb_energy does on the host: the same bits
__global__ void energy_static_finish_kernel(const double *__restrict__ partial, unsigned blocks,
                                            double G, double *__restrict__ ke_pe)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double ke = 0, pe = 0;
    for (unsigned b = 0; b < blocks; ++b) {
        ke += partial[2 * b];
        pe += partial[2 * b + 1];
    }
    ke_pe[0] = 0.5 * ke;
    ke_pe[1] = -G * pe;
}

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
double *g_static_ptr[kMaxDevices]; // device address of g_energy_static_partial, per device
} // namespace

// In host code `g_energy_static_partial` is a stand-in for the symbol, not a device address:
// the address a kernel may use comes from cudaGetSymbolAddress, once per device, with that
// device current.
cudaError_t energy_static(int device, const double *d_pos_xyzm, const double *d_vel_xyz, unsigned n,
                          double G, double *d_ke_pe, cudaStream_t s)
{
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess) return err;
    double *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_ptr[device]) {
            void *p = nullptr;
            err = cudaGetSymbolAddress(&p, g_energy_static_partial);
            if (err != cudaSuccess) return err;
            g_static_ptr[device] = static_cast<double *>(p);
        }
        scratch = g_static_ptr[device];
    }
    const unsigned blocks = energy_blocks(n);
    if (blocks > kEnergyStaticMaxBlocks) return cudaErrorInvalidValue;
    if (blocks) energy_kernel<<<blocks, ETB, 0, s>>>(d_pos_xyzm, d_vel_xyz, n, scratch);
    energy_static_finish_kernel<<<1, 32, 0, s>>>(scratch, blocks, G, d_ke_pe);
    return cudaGetLastError();
}

} // namespace nb
