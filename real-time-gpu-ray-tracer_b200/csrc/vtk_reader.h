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
