#include "vtk_reader.h"
#include "host_parallel.h"

#include <chrono>
#include <climits>
#include <cstdint>
#include <thread>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

// The series file is {"file-series-version":"1.0","files":[{"name":"...","time":t}, ...]}.  A scanning parser
// for exactly that shape: every "name" string followed by its "time" number.
bool VtkSeries::load(const std::string &series_path, std::string &err) {
    std::ifstream f(series_path, std::ios::binary);
    if (!f.is_open()) { err = "cannot open series file " + series_path; return false; }
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string s = ss.str();
    if (s.find("\"files\"") == std::string::npos) { err = "series file has no \"files\" array"; return false; }
    files.clear();
    // every {...} object of the "files" array holds one "name" string and one "time" number, in either order
    size_t pos = s.find('[', s.find("\"files\""));
    if (pos == std::string::npos) { err = "series file: \"files\" is not an array"; return false; }
    const size_t end_of_array = s.find(']', pos);
    for (;;) {
        const size_t open = s.find('{', pos);
        if (open == std::string::npos || (end_of_array != std::string::npos && open > end_of_array)) break;
        const size_t close = s.find('}', open);
        if (close == std::string::npos) { err = "series file: unterminated entry"; return false; }
        const std::string obj = s.substr(open, close - open + 1);
        const size_t k = obj.find("\"name\"");
        if (k == std::string::npos) { err = "series entry without a name"; return false; }
        const size_t c0 = obj.find(':', k);
        const size_t q0 = c0 == std::string::npos ? std::string::npos : obj.find('"', c0 + 1);
        const size_t q1 = q0 == std::string::npos ? std::string::npos : obj.find('"', q0 + 1);
        if (q1 == std::string::npos) { err = "malformed name entry"; return false; }
        const std::string name = obj.substr(q0 + 1, q1 - q0 - 1);
        const size_t t = obj.find("\"time\"");
        if (t == std::string::npos) { err = "name without time"; return false; }
        const size_t colon = obj.find(':', t);
        if (colon == std::string::npos) { err = "malformed time entry"; return false; }
        const float time = std::strtof(obj.c_str() + colon + 1, nullptr);
        files.emplace_back(name, time);
        pos = close + 1;
    }
    const size_t slash = series_path.find_last_of('/');
    dir = slash == std::string::npos ? std::string() : series_path.substr(0, slash + 1);
    if (files.empty()) { err = "series file lists no frames"; return false; }
    return true;
}

// ---- legacy VTK 2.0 BINARY POLYDATA (big-endian), layout in SURVEY.md Appendix B --------------------------------
// Does the job of vtkPolyDataReader in VTKReader::readVTKFile (reference src/Global/VTKReader.cu:16-57): header check,
// POINTS, TRIANGLE_STRIPS, and the presence of the cell arrays "id" and "vel" the reference insists on (:52-57).
namespace {
uint32_t be32(const unsigned char *p) { uint32_t u; memcpy(&u, p, 4); return __builtin_bswap32(u); }
uint64_t be64(const unsigned char *p) { uint64_t u; memcpy(&u, p, 8); return __builtin_bswap64(u); }
// next text line starting at pos (skipping blank lines); returns false at EOF
bool next_line(const std::string &s, size_t &pos, std::string &line) {
    while (pos < s.size()) {
        const size_t e = s.find('\n', pos);
        const size_t end = e == std::string::npos ? s.size() : e;
        line = s.substr(pos, end - pos);
        pos = end + 1;
        while (!line.empty() && (line.back() == ' ' || line.back() == '\r')) line.pop_back();
        if (!line.empty()) return true;
    }
    return false;
}
}

bool VtkFrame::read(const std::string &path, std::string &err) {
    // one read() of the whole file into a buffer that keeps its capacity from frame to frame
    std::string &s = file_;
    {
        FILE *fp = fopen(path.c_str(), "rb");
        if (!fp) { err = "cannot open vtk file " + path; return false; }
        fseek(fp, 0, SEEK_END);
        const long bytes = ftell(fp);
        fseek(fp, 0, SEEK_SET);
        s.resize(bytes > 0 ? (size_t) bytes : 0);
        const size_t got = s.empty() ? 0 : fread(&s[0], 1, s.size(), fp);
        fclose(fp);
        if (got != s.size()) { err = "short read of vtk file " + path; return false; }
    }
    size_t pos = 0;
    std::string line;
    if (!next_line(s, pos, line) || line.find("# vtk DataFile Version") == std::string::npos) { err = "illegal vtk file header in " + path; return false; }
    next_line(s, pos, line);                                       // title
    if (!next_line(s, pos, line) || line.find("BINARY") == std::string::npos) { err = "only BINARY legacy vtk files are supported: " + path; return false; }
    if (!next_line(s, pos, line) || line.find("POLYDATA") == std::string::npos) { err = "DATASET POLYDATA expected in " + path; return false; }
    points.clear();
    strips.clear();
    bool have_points = false, have_strips = false;
    while (next_line(s, pos, line)) {
        std::istringstream ls(line);
        std::string key;
        ls >> key;
        if (key == "POINTS") {
            size_t n = 0;
            std::string type;
            ls >> n >> type;
            const size_t w = type == "double" ? 8 : (type == "float" ? 4 : 0);
            if (!w || pos + n * 3 * w > s.size()) { err = "bad POINTS section in " + path; return false; }
            points.resize(n * 3);
            const unsigned char *p = reinterpret_cast<const unsigned char *>(s.data()) + pos;
            for (size_t i = 0; i < n * 3; i++) {
                if (w == 8) { const uint64_t u = be64(p + 8 * i); double d; memcpy(&d, &u, 8); points[i] = d; }
                else { const uint32_t u = be32(p + 4 * i); float d; memcpy(&d, &u, 4); points[i] = d; }
            }
            pos += n * 3 * w;
            have_points = true;
        } else if (key == "TRIANGLE_STRIPS") {
            size_t cells = 0, size = 0;
            ls >> cells >> size;
            if (pos + size * 4 > s.size()) { err = "bad TRIANGLE_STRIPS section in " + path; return false; }
            const unsigned char *p = reinterpret_cast<const unsigned char *>(s.data()) + pos;
            size_t k = 0;
            strips.resize(cells);
            for (size_t c = 0; c < cells; c++) {
                if (k >= size) { err = "truncated TRIANGLE_STRIPS in " + path; return false; }
                const size_t len = be32(p + 4 * k++);
                if (k + len > size) { err = "truncated TRIANGLE_STRIPS cell in " + path; return false; }
                strips[c].resize(len);
                int *dst = strips[c].data();
                uint32_t worst = 0;
                for (size_t j = 0; j < len; j++) {
                    const uint32_t id = be32(p + 4 * k++);
                    worst = id > worst ? id : worst;
                    dst[j] = (int) id;
                }
                if (len && (size_t) worst * 3 + 2 >= points.size()) { err = "point id out of range in " + path; return false; }
            }
            pos += size * 4;
            have_strips = true;
            break;                                                  // the rest is cell data
        } else if (key == "VERTICES" || key == "LINES" || key == "POLYGONS") {
            err = "found illegal cell type " + key + " (only vtkTriangleStrip cells are accepted, VTKReader.cu:81-85)";
            return false;
        }
    }
    if (!have_points || points.empty() || !have_strips) { err = "no points / strips in " + path; return false; }
    // reference: cell arrays "id" and "vel" must exist (values are not used downstream)
    if (s.find("SCALARS id ", pos) == std::string::npos || s.find("VECTORS vel ", pos) == std::string::npos) { err = "failed to read cell data (id / vel) in " + path; return false; }
    return true;
}

namespace {
// point normals of a frame (restated, parity unpinned — see vtk_reader.h), float64, unit length or zero
void point_normals(const VtkFrame &frame, std::vector<double> &nrm) {
    const std::vector<double> &P = frame.points;
    const size_t n_points = P.size() / 3, n_strips = frame.strips.size();
    nrm.assign(n_points * 3, 0.0);
    // strips that share no point (one particle = its own points, the layout of the series) can be converted concurrently
    bool disjoint = true;
    {
        std::vector<std::pair<int, int>> span(n_strips);
        for (size_t k = 0; k < n_strips; k++) {
            int lo = INT32_MAX, hi = -1;
            for (int id: frame.strips[k]) { lo = std::min(lo, id); hi = std::max(hi, id); }
            span[k] = {lo, hi};
        }
        std::sort(span.begin(), span.end());
        for (size_t k = 1; k < n_strips && disjoint; k++) disjoint = span[k].first > span[k - 1].second;
    }
    // ---- point normals (restated, parity unpinned) ----
    parallel_for(n_strips, disjoint, [&](size_t k) {
        const std::vector<int> &strip = frame.strips[k];
        if (strip.size() < 3) return;
        // centroid of the strip's distinct points, summed in ascending id order
        int lo = INT32_MAX, hi = -1;
        for (int id: strip) { lo = std::min(lo, id); hi = std::max(hi, id); }
        std::vector<unsigned char> seen((size_t) (hi - lo + 1), 0);
        for (int id: strip) seen[id - lo] = 1;
        double c[3] = {0, 0, 0};
        size_t n_uniq = 0;
        for (int id = lo; id <= hi; id++) {
            if (!seen[id - lo]) continue;
            n_uniq++;
            for (int a = 0; a < 3; a++) c[a] += P[3 * id + a];
        }
        for (int a = 0; a < 3; a++) c[a] /= (double) n_uniq;
        for (size_t j = 0; j + 2 < strip.size(); j++) {
            const int i0 = strip[j], i1 = strip[j + 1], i2 = strip[j + 2];
            if (i0 == i1 || i1 == i2 || i0 == i2) continue;
            const double *a = &P[3 * i0], *b = &P[3 * i1], *d = &P[3 * i2];
            const double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {d[0] - a[0], d[1] - a[1], d[2] - a[2]};
            double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
            const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            if (!(len > 0.0)) continue;
            double out = 0.0;
            for (int q = 0; q < 3; q++) { n[q] /= len; out += n[q] * ((a[q] + b[q] + d[q]) / 3.0 - c[q]); }
            const double sgn = out < 0.0 ? -1.0 : 1.0;
            for (int id: {i0, i1, i2}) for (int q = 0; q < 3; q++) nrm[3 * id + q] += sgn * n[q];
        }
    });
    for (size_t i = 0; i < n_points; i++) {
        double *n = &nrm[3 * i];
        const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        if (len > 0.0) { n[0] /= len; n[1] /= len; n[2] /= len; }
    }
}

// per-strip triangle / instance-slot offsets and the instance records (bounds, centroid: VTKReader.cu:97-128)
void strip_layout(const VtkFrame &frame, std::vector<size_t> &first, std::vector<size_t> &slot) {
    const size_t n_strips = frame.strips.size();
    first.assign(n_strips + 1, 0);
    slot.assign(n_strips + 1, 0);
    for (size_t k = 0; k < n_strips; k++) {
        const size_t m = frame.strips[k].size();
        first[k + 1] = first[k] + (m >= 3 ? m - 2 : 0);
        slot[k + 1] = slot[k] + (m >= 3 ? 1 : 0);
    }
}
rt_instance_desc strip_instance(const std::vector<double> &P, const std::vector<int> &strip, size_t first_prim) {
    rt_instance_desc d{};
    d.type = RT_TRIANGLE;
    d.has_bounds = 1;
    d.first_prim = first_prim;
    d.prim_count = strip.size() - 2;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, c[3] = {0, 0, 0};
    for (int id: strip)
        for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], P[3 * id + a]); hi[a] = std::max(hi[a], P[3 * id + a]); c[a] += P[3 * id + a]; }
    for (int a = 0; a < 3; a++) { d.bounds_min[a] = (float) lo[a]; d.bounds_max[a] = (float) hi[a]; d.centroid[a] = (float) (c[a] / (double) strip.size()); }
    return d;
}
const rt_instance_xform PARTICLE_XFORM{{0.0f, 4.0f, 0.0f}, {90.0f, 0.0f, 0.0f}, {3.0f, 3.0f, 3.0f}};     // VTKReader.cu:210-214
} // namespace

void vtk_convert_compact(const VtkFrame &frame, VtkCompact &out, std::vector<rt_instance_desc> &inst, std::vector<rt_instance_xform> &xforms) {
    const std::vector<double> &P = frame.points;
    const size_t n_points = P.size() / 3, n_strips = frame.strips.size();
    std::vector<double> nrm;
#ifdef VTK_PROFILE
    auto T0 = std::chrono::steady_clock::now();
#endif
    point_normals(frame, nrm);
#ifdef VTK_PROFILE
    fprintf(stderr, "point_normals %.2f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - T0).count());
#endif
    out.points.resize(n_points);
    out.normals.resize(n_points);
    for (size_t i = 0; i < n_points; i++) {
        out.points[i] = rt_vec3{(float) P[3 * i], (float) P[3 * i + 1], (float) P[3 * i + 2]};
        out.normals[i] = rt_vec3{(float) nrm[3 * i], (float) nrm[3 * i + 1], (float) nrm[3 * i + 2]};
    }
    std::vector<size_t> first, slot;
    strip_layout(frame, first, slot);
    out.n_tris = first[n_strips];
    out.strip_first_tri.clear();
    out.strip_first_id.clear();
    out.ids.clear();
    inst.assign(slot[n_strips], rt_instance_desc{});
    xforms.assign(slot[n_strips], PARTICLE_XFORM);
    for (size_t k = 0; k < n_strips; k++) {
        const std::vector<int> &strip = frame.strips[k];
        if (strip.size() < 3) continue;
        out.strip_first_tri.push_back((uint32_t) first[k]);
        out.strip_first_id.push_back((uint32_t) out.ids.size());
        out.ids.insert(out.ids.end(), strip.begin(), strip.end());
    }
    out.strip_first_tri.push_back((uint32_t) out.n_tris);          // sentinel
    parallel_for(n_strips, true, [&](size_t k) {
        const std::vector<int> &strip = frame.strips[k];
        if (strip.size() >= 3) inst[slot[k]] = strip_instance(P, strip, first[k]);
    });
}

void vtk_convert(const VtkFrame &frame, std::vector<rt_triangle> &tris, std::vector<rt_instance_desc> &inst,
                 std::vector<rt_instance_xform> &xforms) {
    const std::vector<double> &P = frame.points;
    const size_t n_points = P.size() / 3, n_strips = frame.strips.size();
    std::vector<double> nrm;
    point_normals(frame, nrm);
    // ---- triangles + instances (strips shorter than 3 points give nothing, like the serial loop they replace) ----
    std::vector<size_t> first, slot;
    strip_layout(frame, first, slot);
    tris.resize(first[n_strips]);                 // every element is overwritten below; a caller that keeps the vector pays no page faults
    inst.assign(slot[n_strips], rt_instance_desc{});
    xforms.assign(slot[n_strips], PARTICLE_XFORM);
    // single-precision tables: every point is used by ~30 triangle corners
    std::vector<rt_vec3> Pf(n_points), Nf(n_points);
    for (size_t i = 0; i < n_points; i++) {
        Pf[i] = rt_vec3{(float) P[3 * i], (float) P[3 * i + 1], (float) P[3 * i + 2]};
        Nf[i] = rt_vec3{(float) nrm[3 * i], (float) nrm[3 * i + 1], (float) nrm[3 * i + 2]};
    }
    auto vert = [&](int id) { return Pf[id]; };
    auto norm = [&](int id) { return Nf[id]; };
    parallel_for(n_strips, true, [&](size_t k) {
        const std::vector<int> &strip = frame.strips[k];
        if (strip.size() < 3) return;
        const rt_instance_desc d = strip_instance(P, strip, first[k]);
        rt_triangle *out = tris.data() + first[k];
        for (size_t j = 0; j + 2 < strip.size(); j++) {
            const int i0 = strip[j], i1 = (j & 1) ? strip[j + 2] : strip[j + 1], i2 = (j & 1) ? strip[j + 1] : strip[j + 2];
            rt_triangle t{};
            t.v[0] = vert(i0); t.v[1] = vert(i1); t.v[2] = vert(i2);
            t.n[0] = norm(i0); t.n[1] = norm(i1); t.n[2] = norm(i2);
            t.material_type = RT_METAL; t.material_index = 0; t.has_normals = 1;
            out[j] = t;
        }
        inst[slot[k]] = d;
    });
}
