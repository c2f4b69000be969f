// vtk_reader.h — dependency-free reader for the particle-mesh input of config 3: the `.vtk.series` JSON index
// (reference Renderer::configureVTKFiles, src/Global/Renderer.cu:394-443, which uses nlohmann/json) and the
// legacy-binary VTK POLYDATA frames (reference VTKReader::readVTKFile, src/Global/VTKReader.cu:16-165, which
// uses VTK 9.5).  Host C++ only.
#pragma once

#include <string>
#include <utility>
#include <vector>

struct VtkSeries {
    std::string dir;                                   // directory prefix of the series file (with trailing '/')
    std::vector<std::pair<std::string, float>> files;  // (name relative to dir, time)
    bool load(const std::string &series_path, std::string &err);
};

// One legacy-binary POLYDATA frame: points (as stored, float64) and the triangle-strip cells.
struct VtkFrame {
    std::vector<double> points;                 // 3 per point
    std::vector<std::vector<int>> strips;       // point ids of each TRIANGLE_STRIPS cell (= one particle)
    bool read(const std::string &path, std::string &err);
    std::string file_;                          // the file's bytes (capacity kept between frames)
};

#include "../../include/rtb200.h"

// VTKReader::convertToRendererData (reference src/Global/VTKReader.cu:167-220): every strip of n points becomes n - 2
// triangles (odd ones with vertices 2 and 3 swapped, :184-195), all METAL #0 (:198-201), one instance per particle
// with the cell's bounds and the mean of its strip points (repeats included, float64 sums; :97-128), and the fixed
// transform shift (0,4,0), rotate (90,0,0), scale 3 (:210-214).  Vertex normals restate vtkPolyDataNormals (point
// normals, no splitting, consistent + outward orientation; :60-70) — PARITY UNPINNED, VTK itself cannot run here:
// unit-length sum of the unit normals of the incident non-degenerate triangles, each oriented away from the
// particle's mean point.
void vtk_convert(const VtkFrame &frame, std::vector<rt_triangle> &tris, std::vector<rt_instance_desc> &inst,
                 std::vector<rt_instance_xform> &xforms);

// The same conversion without writing the triangles out on the host: the per-point tables (single precision, as the
// triangle records hold them), the strips' point ids back to back and where each strip's triangles start.  The device
// expands them into triangle records (expand_strips_kernel, kernels_build.cu): 4 MB instead of 53 MB per 500-particle
// frame cross PCIe, and the host writes none of the 629 k records.
struct VtkCompact {
    std::vector<rt_vec3> points, normals;       // per point
    std::vector<int> ids;                       // strips with >= 3 points, concatenated
    std::vector<uint32_t> strip_first_tri;      // per kept strip, plus the total as the last entry
    std::vector<uint32_t> strip_first_id;       // per kept strip: offset of its ids
    size_t n_tris = 0;
};
void vtk_convert_compact(const VtkFrame &frame, VtkCompact &out, std::vector<rt_instance_desc> &inst, std::vector<rt_instance_xform> &xforms);
