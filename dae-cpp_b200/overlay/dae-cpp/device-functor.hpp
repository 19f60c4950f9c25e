/*
 * Device-side functors for dae-cpp on B200 -- an opt-in ADDITION to the dae-cpp API.
 *
 * The reference evaluates the right-hand side and the Jacobian on the host (dae-cpp/solver.hpp:450-464,
 * 499).  For stencil-like problems the user can instead write the equations ONCE, as a trivially copyable
 * functor with a templated call operator -- the device twin of VectorFunctionElements::equations(x, t, i)
 * (dae-cpp/vector-function.hpp:46-72) --
 *
 *     struct MyEquations {
 *         int N; double k;                                            // parameters (differ per ensemble member)
 *         template <class S, class X>
 *         DAECPP_HD S operator()(const X &x, const double t, const int i) const { return k * x[i] - x[i] * x[i] + t; }
 *     };
 *
 * and this header turns it into
 *   - the RHS kernel:        f_i = eq.template operator()<double>(x, t, i)
 *   - the Jacobian kernel:   for every (i, j) of a user-given shape, d f_i / d x_j by forward-mode dual numbers
 *                            seeded in x_j -- what JacobianMatrixShape does with autodiff
 *                            (dae-cpp/jacobian-matrix.hpp:66-92), values in the order of the shape list
 * behind the two callbacks of `daecpp_b200_device_functors` (include/daecpp_b200.h).  State, right-hand side
 * and Jacobian values then never leave HBM.
 *
 * Compile the translation unit that instantiates DeviceFunctors<Eq> with nvcc for sm_100a; add -fmad=false
 * when the device values must be bit-identical to the host instantiation of the same functor (host
 * compilers do not contract a*b+c on x86-64 by default, nvcc does).  Compiled as plain C++ (no __CUDACC__)
 * the same classes run the functor in host loops over HOST pointers: that mode exists for the CPU tests of
 * the host-side logic (tests/emu) and is not a product path.
 */
#ifndef DAECPP_DEVICE_FUNCTOR_H
#define DAECPP_DEVICE_FUNCTOR_H

#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <type_traits>
#include <utility>
#include <vector>

#include <daecpp_b200.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define DAECPP_HD __host__ __device__ __forceinline__
#else
#define DAECPP_HD inline
#endif

namespace daecpp
{
namespace device
{

// First-order dual number: value and one directional derivative.
struct dual
{
    double v, d;
    DAECPP_HD dual() : v(0.0), d(0.0) {}
    DAECPP_HD dual(double v_) : v(v_), d(0.0) {}
    DAECPP_HD dual(double v_, double d_) : v(v_), d(d_) {}
};
DAECPP_HD dual operator+(const dual &a, const dual &b) { return dual(a.v + b.v, a.d + b.d); }
DAECPP_HD dual operator-(const dual &a, const dual &b) { return dual(a.v - b.v, a.d - b.d); }
DAECPP_HD dual operator*(const dual &a, const dual &b) { return dual(a.v * b.v, a.d * b.v + a.v * b.d); }
DAECPP_HD dual operator/(const dual &a, const dual &b)
{
    const double q = a.v / b.v;
    return dual(q, (a.d - q * b.d) / b.v);
}
DAECPP_HD dual operator-(const dual &a) { return dual(-a.v, -a.d); }
DAECPP_HD dual operator+(const dual &a) { return a; }
DAECPP_HD dual operator+(const dual &a, double b) { return dual(a.v + b, a.d); }
DAECPP_HD dual operator+(double a, const dual &b) { return dual(a + b.v, b.d); }
DAECPP_HD dual operator-(const dual &a, double b) { return dual(a.v - b, a.d); }
DAECPP_HD dual operator-(double a, const dual &b) { return dual(a - b.v, -b.d); }
DAECPP_HD dual operator*(const dual &a, double b) { return dual(a.v * b, a.d * b); }
DAECPP_HD dual operator*(double a, const dual &b) { return dual(a * b.v, a * b.d); }
DAECPP_HD dual operator/(const dual &a, double b) { return dual(a.v / b, a.d / b); }
DAECPP_HD dual operator/(double a, const dual &b)
{
    const double q = a / b.v;
    return dual(q, -q * b.d / b.v);
}
DAECPP_HD dual &operator+=(dual &a, const dual &b) { a = a + b; return a; }
DAECPP_HD dual &operator-=(dual &a, const dual &b) { a = a - b; return a; }
DAECPP_HD dual &operator*=(dual &a, const dual &b) { a = a * b; return a; }
DAECPP_HD dual &operator/=(dual &a, const dual &b) { a = a / b; return a; }
DAECPP_HD dual exp(const dual &a) { const double e = ::exp(a.v); return dual(e, e * a.d); }
DAECPP_HD dual log(const dual &a) { return dual(::log(a.v), a.d / a.v); }
DAECPP_HD dual sqrt(const dual &a) { const double s = ::sqrt(a.v); return dual(s, 0.5 * a.d / s); }
DAECPP_HD dual sin(const dual &a) { return dual(::sin(a.v), ::cos(a.v) * a.d); }
DAECPP_HD dual cos(const dual &a) { return dual(::cos(a.v), -::sin(a.v) * a.d); }
DAECPP_HD dual tanh(const dual &a) { const double t = ::tanh(a.v); return dual(t, (1.0 - t * t) * a.d); }
DAECPP_HD dual pow(const dual &a, double p) { return dual(::pow(a.v, p), p * ::pow(a.v, p - 1.0) * a.d); }
// the same names for plain doubles, so that a functor can call them through ADL-free `using`
DAECPP_HD double exp(double a) { return ::exp(a); }
DAECPP_HD double log(double a) { return ::log(a); }
DAECPP_HD double sqrt(double a) { return ::sqrt(a); }
DAECPP_HD double sin(double a) { return ::sin(a); }
DAECPP_HD double cos(double a) { return ::cos(a); }
DAECPP_HD double tanh(double a) { return ::tanh(a); }
DAECPP_HD double pow(double a, double p) { return ::pow(a, p); }

// what a functor sees as `x`: the lane's state, as doubles or as duals seeded in one column
struct StateView
{
    const double *x;
    DAECPP_HD double operator[](const int k) const { return x[k]; }
};
struct SeededView
{
    const double *x;
    int32_t j;
    DAECPP_HD dual operator[](const int k) const { return dual(x[k], k == j ? 1.0 : 0.0); }
};

// The fixed list of non-zero Jacobian elements, in the order the triplets are handed to the solver: the
// add_element interface of JacobianMatrixShape (dae-cpp/jacobian-matrix.hpp:94-112).
class Shape
{
    std::vector<int32_t> m_i, m_j;

public:
    void add_element(const int32_t i, const int32_t j) { m_i.push_back(i); m_j.push_back(j); }
    void add_element(const int32_t i, const std::vector<int32_t> &cols)
    {
        for (const int32_t j : cols) add_element(i, j);
    }
    void reserve(const std::size_t count) { m_i.reserve(count); m_j.reserve(count); }
    void clear() { m_i.clear(); m_j.clear(); }
    std::size_t size() const { return m_i.size(); }
    const std::vector<int32_t> &rows() const { return m_i; }
    const std::vector<int32_t> &cols() const { return m_j; }
};

// Optional: a functor may also provide its Jacobian elements in closed form,
//     template <class X> DAECPP_HD double jacobian(const X &x, const double t, const int i, const int j) const;
// (the device twin of a hand-written JacobianMatrix functor, dae-cpp/jacobian-matrix.hpp:30-43).  When present it is
// used for the shape's elements instead of differentiating `operator()`.
namespace detail
{
template <class Eq, class = void> struct has_jacobian : std::false_type {};
template <class Eq>
struct has_jacobian<Eq, std::void_t<decltype(std::declval<const Eq &>().jacobian(std::declval<const StateView &>(), 0.0, 0, 0))>> : std::true_type {};
template <class Eq> DAECPP_HD double jacobian_element(const Eq &eq, const double *x, const double t, const int32_t i, const int32_t j)
{
    if constexpr (has_jacobian<Eq>::value) return eq.jacobian(StateView{x}, t, i, j);
    else return eq.template operator()<dual>(SeededView{x, j}, t, i).d;
}
} // namespace detail

#if defined(__CUDACC__)
namespace detail
{
// The functor kernels sit between the library's own kernels on the stream and are launched like them with programmatic
// stream serialisation: a kernel lets its successor start launching at once and waits for its predecessor's memory
// before it reads anything (every path through the kernel passes the wait: the chain of dependencies stays transitive).
__device__ __forceinline__ void pdl_entry()
{
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 900
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
template <class... P, class... A> inline cudaError_t launch_pdl(cudaStream_t st, void (*kernel)(P...), dim3 grid, int threads, A... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3((unsigned)threads, 1, 1);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}
template <class Eq>
__global__ void k_rhs(const Eq *__restrict__ eq, const daecpp_b200_functor_args a)
{
    pdl_entry();
    const int lane = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    if (a.mode[lane] < a.min_mode) return; // (also lanes that halted earlier in a multi-iteration round)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const StateView x{a.x + (int64_t)lane * a.x_stride};
    a.out[(int64_t)lane * a.out_stride + i] = eq[lane].template operator()<double>(x, a.t[lane], i);
}
template <class Eq>
__global__ void k_jacobian(const Eq *__restrict__ eq, const int32_t *__restrict__ ti, const int32_t *__restrict__ tj,
                           const int64_t nT, const daecpp_b200_functor_args a)
{
    pdl_entry();
    const int lane = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    if (a.mode[lane] < a.min_mode) return; // (also lanes that halted earlier in a multi-iteration round)
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nT) return;
    a.out[(int64_t)lane * a.out_stride + k] = jacobian_element(eq[lane], a.x + (int64_t)lane * a.x_stride, a.t[lane], ti[k], tj[k]);
}
} // namespace detail
#endif

/*
 * Owns the per-lane functor objects and the Jacobian shape (in device memory when compiled by nvcc) and
 * provides the callbacks of `daecpp_b200_device_functors`.  One object serves one batch context; it must
 * outlive it.
 */
template <class Eq>
class DeviceFunctors
{
    Eq *m_eq{nullptr};
    int32_t *m_i{nullptr}, *m_j{nullptr};
    int64_t m_nT{0};
    int32_t m_lanes{0};

    static void fail(const char *what) { throw std::runtime_error(std::string("daecpp::device::DeviceFunctors: ") + what); }

#if defined(__CUDACC__)
    template <class T> static T *to_device(const T *h, std::size_t count)
    {
        T *d = nullptr;
        if (cudaMalloc((void **)&d, sizeof(T) * (count ? count : 1)) != cudaSuccess) fail("cudaMalloc failed");
        if (count && cudaMemcpy(d, h, sizeof(T) * count, cudaMemcpyHostToDevice) != cudaSuccess) fail("cudaMemcpy failed");
        return d;
    }
    template <class T> static void release(T *p) { if (p) cudaFree(p); }
    static int rhs_cb(void *user, const daecpp_b200_functor_args *a)
    {
        const DeviceFunctors *self = static_cast<const DeviceFunctors *>(user);
        if (a->lanes != self->m_lanes) return 1;
        if (a->lane_list && a->n_list == 0) return 0;
        const dim3 grid((unsigned)((a->n + 255) / 256), (unsigned)(a->lane_list ? a->n_list : a->lanes));
        const cudaError_t e = detail::launch_pdl(static_cast<cudaStream_t>(a->stream), detail::k_rhs<Eq>, grid, 256, self->m_eq, *a);
        return (e == cudaSuccess && cudaGetLastError() == cudaSuccess) ? 0 : 2;
    }
    static int jac_cb(void *user, const daecpp_b200_functor_args *a)
    {
        const DeviceFunctors *self = static_cast<const DeviceFunctors *>(user);
        if (a->lanes != self->m_lanes) return 1;
        if (a->lane_list && a->n_list == 0) return 0;
        const dim3 grid((unsigned)((self->m_nT + 255) / 256), (unsigned)(a->lane_list ? a->n_list : a->lanes));
        const cudaError_t e = detail::launch_pdl(static_cast<cudaStream_t>(a->stream), detail::k_jacobian<Eq>, grid, 256, self->m_eq, self->m_i, self->m_j, self->m_nT, *a);
        return (e == cudaSuccess && cudaGetLastError() == cudaSuccess) ? 0 : 2;
    }
#else
    template <class T> static T *to_device(const T *h, std::size_t count)
    {
        T *d = static_cast<T *>(::operator new(sizeof(T) * (count ? count : 1)));
        for (std::size_t k = 0; k < count; ++k) new (d + k) T(h[k]);
        return d;
    }
    template <class T> static void release(T *p) { ::operator delete(p); }
    static int rhs_cb(void *user, const daecpp_b200_functor_args *a)
    {
        const DeviceFunctors *self = static_cast<const DeviceFunctors *>(user);
        if (a->lanes != self->m_lanes) return 1;
        for (int32_t q = 0; q < (a->lane_list ? a->n_list : a->lanes); ++q)
        {
            const int32_t lane = a->lane_list ? a->lane_list[q] : q;
            if (a->mode[lane] < a->min_mode) continue;
            const StateView x{a->x + (int64_t)lane * a->x_stride};
            for (int32_t i = 0; i < a->n; ++i)
                a->out[(int64_t)lane * a->out_stride + i] = self->m_eq[lane].template operator()<double>(x, a->t[lane], i);
        }
        return 0;
    }
    static int jac_cb(void *user, const daecpp_b200_functor_args *a)
    {
        const DeviceFunctors *self = static_cast<const DeviceFunctors *>(user);
        if (a->lanes != self->m_lanes) return 1;
        for (int32_t q = 0; q < (a->lane_list ? a->n_list : a->lanes); ++q)
        {
            const int32_t lane = a->lane_list ? a->lane_list[q] : q;
            if (a->mode[lane] < a->min_mode) continue;
            for (int64_t k = 0; k < self->m_nT; ++k)
            {
                a->out[(int64_t)lane * a->out_stride + k] =
                    detail::jacobian_element(self->m_eq[lane], a->x + (int64_t)lane * a->x_stride, a->t[lane], self->m_i[k], self->m_j[k]);
            }
        }
        return 0;
    }
#endif

public:
    // one functor object per lane (the members of a parameter sweep differ in their parameters only)
    DeviceFunctors(const std::vector<Eq> &per_lane, const Shape &shape)
        : m_nT((int64_t)shape.size()), m_lanes((int32_t)per_lane.size())
    {
        if (per_lane.empty()) fail("no lanes");
        m_eq = to_device(per_lane.data(), per_lane.size());
        m_i = to_device(shape.rows().data(), shape.size());
        m_j = to_device(shape.cols().data(), shape.size());
    }
    DeviceFunctors(const DeviceFunctors &) = delete;
    DeviceFunctors &operator=(const DeviceFunctors &) = delete;
    ~DeviceFunctors()
    {
        release(m_eq);
        release(m_i);
        release(m_j);
    }
    int32_t lanes() const { return m_lanes; }
    daecpp_b200_device_functors binding() { return daecpp_b200_device_functors{this, &DeviceFunctors::rhs_cb, &DeviceFunctors::jac_cb}; }
};

} // namespace device
} // namespace daecpp

#endif // DAECPP_DEVICE_FUNCTOR_H
