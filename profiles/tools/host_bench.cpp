// host_bench.cpp — times the host half of one animation frame (file -> triangles -> device-layout records) without a GPU:
//   g++ -O2 -std=c++17 -I/usr/local/cuda/include -Iinclude -Ireal-time-gpu-ray-tracer_b200/csrc profiles/tools/host_bench.cpp \
//       real-time-gpu-ray-tracer_b200/csrc/vtk_reader.cpp real-time-gpu-ray-tracer_b200/csrc/host_scene.cpp -lpthread -o host_bench
//   ./host_bench frame.vtk [repeats]
#include "vtk_reader.h"
#include "host_scene.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: host_bench frame.vtk [repeats]\n"); return 2; }
    const int reps = argc > 2 ? atoi(argv[2]) : 10;
    VtkFrame vf;
    std::vector<rt_triangle> tris;
    std::vector<rt_instance_desc> inst;
    std::vector<rt_instance_xform> xf;
    HostScene hs;
    using clk = std::chrono::steady_clock;
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    double best[4] = {1e30, 1e30, 1e30, 1e30};
    VtkCompact vc;
    for (int r = 0; r < reps; r++) {
        std::string err;
        const auto t0 = clk::now();
        if (!vf.read(argv[1], err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        const auto t1 = clk::now();
        vtk_convert(vf, tris, inst, xf);
        const auto t2 = clk::now();
        const std::string e = hs.commit_geometry(nullptr, 0, nullptr, 0, tris.data(), tris.size());
        const auto t3 = clk::now();
        if (!e.empty()) { fprintf(stderr, "%s\n", e.c_str()); return 1; }
        vtk_convert_compact(vf, vc, inst, xf);                // what the read-ahead thread does when the device expands the strips
        const auto t4 = clk::now();
        best[0] = std::min(best[0], ms(t0, t1)); best[1] = std::min(best[1], ms(t1, t2)); best[2] = std::min(best[2], ms(t2, t3));
        best[3] = std::min(best[3], ms(t3, t4));
    }
    unsigned long long h = 1469598103934665603ull;          // FNV-1a over the outputs: optimisations must not change a bit
    auto mix = [&](const void *p, size_t n) { const unsigned char *b = (const unsigned char *) p; for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; } };
    mix(hs.tri_verts.data(), hs.tri_verts.size() * 4);
    mix(hs.tri_shade.data(), hs.tri_shade.size() * sizeof(hs.tri_shade[0]));
    mix(inst.data(), inst.size() * sizeof(inst[0]));
    printf("triangles %zu instances %zu: read %.2f ms, convert %.2f ms, commit %.2f ms, checksum %016llx | compact convert %.2f ms (%.1f MB to upload instead of %.1f)\n",
           tris.size(), inst.size(), best[0], best[1], best[2], h, best[3],
           (vc.points.size() * 24 + vc.ids.size() * 4) / 1e6, tris.size() * 84 / 1e6);
    return 0;
}
