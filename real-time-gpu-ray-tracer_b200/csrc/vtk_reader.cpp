#include "vtk_reader.h"

#include <cstdlib>
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
    size_t pos = s.find("\"files\"");
    for (;;) {
        const size_t k = s.find("\"name\"", pos);
        if (k == std::string::npos) break;
        const size_t q0 = s.find('"', s.find(':', k) + 1);
        const size_t q1 = s.find('"', q0 + 1);
        if (q0 == std::string::npos || q1 == std::string::npos) { err = "malformed name entry"; return false; }
        const std::string name = s.substr(q0 + 1, q1 - q0 - 1);
        const size_t t = s.find("\"time\"", q1);
        if (t == std::string::npos) { err = "name without time"; return false; }
        const size_t colon = s.find(':', t);
        const float time = std::strtof(s.c_str() + colon + 1, nullptr);
        files.emplace_back(name, time);
        pos = colon;
    }
    const size_t slash = series_path.find_last_of('/');
    dir = slash == std::string::npos ? std::string() : series_path.substr(0, slash + 1);
    if (files.empty()) { err = "series file lists no frames"; return false; }
    return true;
}
