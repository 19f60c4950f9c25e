// Device-functor statements of the BASELINE problems (the same equations traj_dump.cpp hands to the host
// API, written once as templated functors for dae-cpp/device-functor.hpp).  Test / bench drivers only.
#pragma once
#include <dae-cpp/device-functor.hpp>

#include <cmath>
#include <vector>

namespace problems
{

// C2 / C5 member: examples/perovskite_model (ion-migration Poisson-Nernst-Planck, singular mass matrix);
// x = [P_0..P_{N-1}, Phi_0..Phi_{N-1}]   (examples/perovskite_model/perovskite.cpp:60-135)
struct PerovskiteEq
{
    int N;
    double ih2, il2; // 1/h^2, 1/lambda^2
    static PerovskiteEq make(int N, double L, double lambda)
    {
        const double h = L / double(N - 1);
        return PerovskiteEq{N, 1.0 / (h * h), 1.0 / (lambda * lambda)};
    }
    template <class S, class X>
    DAECPP_HD S operator()(const X &x, const double t, const int k) const
    {
        if (k == 0) return (x[1] - x[0] + 0.5 * (x[1] + x[0]) * (x[N + 1] - x[N])) * ih2;
        if (k < N - 1)
            return (x[k + 1] - 2.0 * x[k] + x[k - 1] +
                    0.5 * ((x[k + 1] + x[k]) * (x[N + k + 1] - x[N + k]) - (x[k] + x[k - 1]) * (x[N + k] - x[N + k - 1]))) * ih2;
        if (k == N - 1) return -(x[N - 1] - x[N - 2] + 0.5 * (x[N - 1] + x[N - 2]) * (x[2 * N - 1] - x[2 * N - 2])) * ih2;
        if (k == N) return x[N] + t;
        if (k < 2 * N - 1) return (x[k + 1] - 2.0 * x[k] + x[k - 1]) * ih2 - (1.0 - x[k - N]) * il2;
        return x[2 * N - 1] - t;
    }
};
// The same member with its Jacobian in closed form: the expressions of the example's hand-written Jacobian functor
// (examples/perovskite_model/perovskite.cpp:132-186, restated in hostcpp/traj_dump.cpp: PerovskiteJac), so that the values
// handed to the solver are those of the host functor bit for bit.
struct PerovskiteEqAnalytic : PerovskiteEq
{
    explicit PerovskiteEqAnalytic(const PerovskiteEq &e) : PerovskiteEq(e) {}
    template <class X>
    DAECPP_HD double jacobian(const X &x, const double, const int r, const int c) const
    {
        if (r == 0)
        {
            const double dphi = x[N + 1] - x[N], ps = x[0] + x[1];
            if (c == 0) return (-1.0 + 0.5 * dphi) * ih2;
            if (c == 1) return (1.0 + 0.5 * dphi) * ih2;
            if (c == N) return -0.5 * ps * ih2;
            return 0.5 * ps * ih2;
        }
        if (r < N - 1)
        {
            if (c == r - 1) return (1.0 - 0.5 * (x[N + r] - x[N + r - 1])) * ih2;
            if (c == r) return (-2.0 + 0.5 * (x[N + r + 1] - 2.0 * x[N + r] + x[N + r - 1])) * ih2;
            if (c == r + 1) return (1.0 + 0.5 * (x[N + r + 1] - x[N + r])) * ih2;
            if (c == N + r - 1) return 0.5 * (x[r] + x[r - 1]) * ih2;
            if (c == N + r) return -0.5 * (x[r + 1] + 2.0 * x[r] + x[r - 1]) * ih2;
            return 0.5 * (x[r + 1] + x[r]) * ih2;
        }
        if (r == N - 1)
        {
            const double dphi = x[2 * N - 1] - x[2 * N - 2], ps = x[N - 1] + x[N - 2];
            if (c == r - 1) return (1.0 - 0.5 * dphi) * ih2;
            if (c == r) return (-1.0 - 0.5 * dphi) * ih2;
            if (c == N + r - 1) return 0.5 * ps * ih2;
            return -0.5 * ps * ih2;
        }
        if (r == N || r == 2 * N - 1) return 1.0;
        if (c == r - N) return il2;
        if (c == r) return -2.0 * ih2;
        return ih2;
    }
};
inline daecpp::device::Shape perovskite_shape(int N)
{
    daecpp::device::Shape s;
    s.reserve(10 * (std::size_t)N);
    for (int r = 0; r < N; ++r)
    {
        for (int c : {r - 1, r, r + 1}) if (c >= 0 && c < N) s.add_element(r, c);
        for (int c : {r - 1, r, r + 1}) if (c >= 0 && c < N) s.add_element(r, N + c);
    }
    s.add_element(N, N);
    for (int r = N + 1; r < 2 * N - 1; ++r) s.add_element(r, std::vector<int32_t>{r - N, r - 1, r, r + 1});
    s.add_element(2 * N - 1, 2 * N - 1);
    return s;
}

// C3: 1-D flame front u_t = D u_xx + u^2 - u^3, zero-flux ends
struct Flame1DEq
{
    int n;
    double c; // D / h^2
    static Flame1DEq make(int n, double D = 1e-3)
    {
        const double h = 1.0 / double(n - 1);
        return Flame1DEq{n, D * (1.0 / (h * h))};
    }
    template <class S, class X>
    DAECPP_HD S operator()(const X &u, const double, const int k) const
    {
        const S react = u[k] * u[k] - u[k] * u[k] * u[k];
        if (k == 0) return c * (u[1] - u[0]) + react;
        if (k == n - 1) return c * (u[n - 2] - u[n - 1]) + react;
        return c * (u[k + 1] - 2.0 * u[k] + u[k - 1]) + react;
    }
};
inline daecpp::device::Shape flame1d_shape(int n)
{
    daecpp::device::Shape s;
    s.reserve(3 * (std::size_t)n);
    for (int k = 0; k < n; ++k)
        for (int c : {k - 1, k, k + 1}) if (c >= 0 && c < n) s.add_element(k, c);
    return s;
}
inline std::vector<double> flame1d_initial(int n, double delta = 1e-4)
{
    std::vector<double> u0((std::size_t)n);
    for (int k = 0; k < n; ++k)
    {
        const double x = double(k) / double(n - 1);
        u0[(std::size_t)k] = delta + (1.0 - delta) * 0.5 * (1.0 - std::tanh((x - 0.1) / 0.02));
    }
    return u0;
}

} // namespace problems
