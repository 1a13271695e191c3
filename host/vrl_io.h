// vrl_io.h -- file formats either side of the path, host side, no Mitsuba dependency:
//   * Mitsuba binary grid volumes `.vol` (src/textures/volume_data.h:42-107: "VOL", u8 version 3, i32 type 1 = fp32,
//     i32 nx ny nz, i32 channels, 6 x f32 bbox, data with x fastest and channels innermost, grid3d.cpp:291);
//   * Wavefront OBJ meshes as the fixtures use them (v / f records, polygons fanned, negative indices; normals and
//     texture coordinates are parsed past -- the gather only needs positions, src/shapes/obj.cpp);
//   * the developed film: HDRFilm::develop divides XYZ by W and converts XYZ -> linear sRGB (hdrfilm.cpp:259-262,
//     include/mitsuba/core/spectrum.h:229-235); written as a PFM image or as an uncompressed float OpenEXR file.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace vrlhost {

struct VolumeGrid {
    int32_t nx = 0, ny = 0, nz = 0, channels = 1;
    float bbox_min[3] = { 0, 0, 0 }, bbox_max[3] = { 1, 1, 1 };
    std::vector<float> data;        // ((z * ny + y) * nx + x) * channels + c
    float  max = 0.f;               // VolumeMetadata::max  (the majorant before `scale`)
    double mean = 0.0;              // VolumeMetadata::mean
};
VolumeGrid read_vol(const std::string &path);                    // throws std::runtime_error with the reference's messages
void write_vol(const std::string &path, const VolumeGrid &g);

// triangles as 9 floats each (p0 p1 p2), `to_world` = optional row-major 4x4 applied to the positions
std::vector<float> read_obj(const std::string &path, const float *to_world = nullptr);

// XYZAW (5 floats / pixel) -> linear sRGB (3 floats / pixel); pixels with W = 0 stay black
std::vector<float> develop_rgb(const std::vector<float> &xyzaw, int width, int height);
void write_pfm(const std::string &path, const std::vector<float> &rgb, int width, int height);   // top row first in `rgb`
// The developed film as the reference's HDRFilm writes it (hdrfilm.cpp:254-349, pixel_format "rgba"): an OpenEXR scan-line
// file with R, G, B = xyz_to_srgb(XYZ / W) and A = A / W.  Written uncompressed with 32-bit float channels (the reference's
// default is half floats + ZIP through the OpenEXR library, which is not in this image; any EXR reader takes both).
void write_exr(const std::string &path, const std::vector<float> &xyzaw, int width, int height, bool half = false);   // half: 16-bit channels (component_format float16, the film's default)
// `out` by extension: .exr -> write_exr, .pfm -> develop_rgb + write_pfm, anything else -> the raw XYZAW floats
void write_film(const std::string &out, const std::vector<float> &xyzaw, int width, int height, bool half = false);

}  // namespace vrlhost
