// host_parallel.h — fn(k) for k in [0, n) on up to 8 host threads (contiguous chunks); serial when n is small.
#pragma once

#include <algorithm>
#include <cstddef>
#include <thread>
#include <vector>

template<class F>
static void parallel_for(size_t n, bool parallel, F fn, size_t min_per_thread = 1) {
    const unsigned hw = std::thread::hardware_concurrency();
    size_t T = parallel ? std::min<size_t>(std::min<size_t>(hw ? hw : 1u, 8u), n / std::max<size_t>(1, min_per_thread)) : 1;
    if (T <= 1) { for (size_t k = 0; k < n; k++) fn(k); return; }
    std::vector<std::thread> th;
    for (size_t t = 0; t < T; t++)
        th.emplace_back([=]() { for (size_t k = n * t / T; k < n * (t + 1) / T; k++) fn(k); });
    for (std::thread &x: th) x.join();
}
