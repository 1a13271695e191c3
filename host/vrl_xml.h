// vrl_xml.h -- the subset of Mitsuba 2's scene XML that the reference's VRL fixture scenes use
// (data/scenes/cbox/{cbox-volume-vrl, cbox-volume-nl-vrl, cbox-nonlinear-homo-vrl, mirage-vrl, cbox-heterogeneous-vrl}.xml
// and the files they include), loaded into the host-side mirror types of vrl_integrator.h.  It replaces, for this
// path only, src/libcore/xml.cpp (parse, <default>/$param substitution :618-649, <include> :651-690, typed property
// tags, <transform> operations :806-905, <ref>) plus the constructors of the plugins the fixtures instantiate:
//   sensor   perspective (src/sensors/perspective.cpp:  fov, fov_axis, near_clip, far_clip, to_world) with its nested
//            independent sampler (sample_count, seed) and hdrfilm (width, height)
//   shape    obj, rectangle, cube, sphere (src/shapes/*.cpp; the sphere is tessellated -- the reference intersects it
//            analytically -- and reported in `warnings`), bsdf by nesting or <ref>, interior / exterior medium, area emitter
//   medium   homogeneous, heterogeneous (sigma_t = gridvolume read with read_vol), nonlinear (graded IOR, bbox = the
//            shapes it is interior to, nonlinear.cpp:62-89); phase isotropic / hg
//   integrator   type "vrl": every child becomes a property of vrlhost::Properties, so the unqueried-property and
//            wrong-type errors are those of VRLIntegrator's constructor
// What the gather asks of a BSDF is one bit (BSDFFlags::Null or not, vrl_struct.h:529); the camera path and the light
// tracer additionally read the diffuse reflectance, and `conductor` / `dielectric` load as the delta classes of the
// specular chains (VRLB_SURF_BSDF_MIRROR / _DIELECTRIC).  Anything else is reduced as listed in `warnings`.
// Scene versions before 2.0 get the reference's in-memory upgrade (xml.cpp:350-392: camelCase -> underscore_case names,
// lookAt) and, for the 0.x scenes of data/FBD15ALVRL_scenes, the schema translation of SURVEY.md 8c (sigma_s/sigma_a ->
// sigma_t/albedo, medium boundary without BSDF -> null BSDF, the other implementation's integrator properties dropped);
// every step is listed in `warnings`.
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "vrl_integrator.h"

namespace vrlhost {

struct LoadedScene {
    HostScene   scene;
    HostSensor  sensor;
    Properties  integrator{ "vrl" };
    std::string integrator_type;                 // "" when the file has no <integrator>
    bool film_half = true;                       // hdrfilm component_format: float16 (the default, hdrfilm.cpp:108-109) or float32
    std::vector<std::string> warnings;
    std::vector<std::string> media_ids;          // XML id of media[i]
    std::vector<std::unique_ptr<std::vector<float>>> grids;   // storage behind vrlb_medium::grid pointers
};

// `params` are -Dname=value style overrides of <default> tags.  Throws std::runtime_error with file / tag context.
LoadedScene load_scene_xml(const std::string &path, const std::map<std::string, std::string> &params = {});

}  // namespace vrlhost
