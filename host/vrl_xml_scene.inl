// vrl_xml_scene.inl -- part 2 of vrl_xml.cpp: plugin tags -> HostScene / HostSensor / Properties.
namespace vrlhost {
namespace {

struct Builder {
    LoadedScene out;
    std::string base_dir;
    std::map<std::string, const Node *> by_id;          // id -> bsdf / medium / emitter node
    std::map<std::string, int> medium_index, emitter_index;
    std::vector<std::array<double, 6>> nl_boxes;        // per medium: AABB of the shapes it is interior to
    std::vector<bool> nl_box_set;

    void warn(const std::string &w) { if (std::find(out.warnings.begin(), out.warnings.end(), w) == out.warnings.end()) out.warnings.push_back(w); }

    void index_ids(const Node &n) {
        for (const Node &c : n.children) {
            if (c.name != "ref" && c.attr("id")) {
                if (by_id.count(*c.attr("id"))) fail("Error while loading \"" + c.file + "\": <" + c.name + "> has duplicate id \"" + *c.attr("id") + "\"");
                by_id[*c.attr("id")] = &c;
            }
            index_ids(c);
        }
    }
    const Node &deref(const Node &ref) {
        auto it = by_id.find(ref.get("id"));
        if (it == by_id.end()) fail("Error while loading \"" + ref.file + "\": reference to unknown object \"" + ref.get("id") + "\"");
        return *it->second;
    }

    // ---- media ----
    int medium(const Node &m) {
        const std::string id = m.get("id", "");
        if (!id.empty() && medium_index.count(id)) return medium_index[id];
        const std::string type = m.get("type");
        vrlb_medium d{};
        d.scale = (float) prop_float(m, "scale", 1.0);
        d.phase = VRLB_PHASE_ISOTROPIC; d.g = 0.f;
        if (const Node *ph = child_named(m, "phase", nullptr)) {
            const std::string pt = ph->get("type");
            if (pt == "isotropic") {}
            else if (pt == "hg") { d.phase = VRLB_PHASE_HG; d.g = (float) prop_float(*ph, "g", 0.8); }       // hg.cpp:31
            else fail("Error while loading \"" + m.file + "\": phase function \"" + pt + "\" is not on the VRL path");
        }
        for (int k = 0; k < 3; ++k) { d.albedo[k] = 0.75f; d.sigma_t[k] = 1.f; }                             // homogeneous.cpp:23-24
        if (const Node *al = prop(m, "albedo")) {
            // a constant: <rgb|spectrum|float name="albedo">, or <volume name="albedo" type="constvolume"> with its `color` / `value`
            bool ok = prop_rgb(m, "albedo", d.albedo);
            if (!ok && al->name == "volume" && al->get("type", "") == "constvolume") ok = prop_rgb(*al, "color", d.albedo) || prop_rgb(*al, "value", d.albedo);
            if (!ok) fail("Error while loading \"" + m.file + "\": medium \"" + id + "\": only a constant albedo is supported");
        }
        if (type == "homogeneous" || type == "nonlinear") {
            d.kind = type == "homogeneous" ? VRLB_MEDIUM_HOMOGENEOUS : VRLB_MEDIUM_NONLINEAR;
            if (prop(m, "sigma_t") && !prop_rgb(m, "sigma_t", d.sigma_t))
                fail("Error while loading \"" + m.file + "\": medium \"" + id + "\": sigma_t must be a constant for type \"" + type + "\"");
            if (type == "nonlinear") {                                                                        // nonlinear.cpp:46-57
                d.top_ior = (float) prop_float(m, "top_ior", 0.7); d.bottom_ior = (float) prop_float(m, "bottom_ior", 1.0);
                d.ior_method = (int32_t) prop_int(m, "ior_method", 0); d.from_bottom = prop_bool(m, "from_bottom", true) ? 1 : 0;
                d.nl_res[0] = (int32_t) prop_int(m, "res_x", 4); d.nl_res[1] = (int32_t) prop_int(m, "res_y", 4); d.nl_res[2] = (int32_t) prop_int(m, "res_z", 4);
                if (d.ior_method != 0) fail("Error while loading \"" + m.file + "\": ior_method " + std::to_string(d.ior_method) + " is not on the VRL path (only 0, linear in height)");
            }
        } else if (type == "heterogeneous") {
            d.kind = VRLB_MEDIUM_GRID;
            const Node *vol = prop(m, "sigma_t");
            if (!vol || vol->name != "volume" || vol->get("type") != "gridvolume")
                fail("Error while loading \"" + m.file + "\": heterogeneous medium \"" + id + "\" needs <volume name=\"sigma_t\" type=\"gridvolume\">");
            VolumeGrid g = read_vol(resolve(base_dir, prop_string(*vol, "filename", "")));
            if (g.channels != 1) fail("Error while loading \"" + m.file + "\": sigma_t grid must have one channel");
            // grid3d.cpp:63-75: with use_grid_bbox the file's bbox maps onto [0,1]^3 first; to_world then places the unit cube
            Mat4 tw = to_world_of(*vol);
            double lo[3] = { 0, 0, 0 }, hi[3] = { 1, 1, 1 };
            if (prop_bool(*vol, "use_grid_bbox", false)) for (int k = 0; k < 3; ++k) { lo[k] = g.bbox_min[k]; hi[k] = g.bbox_max[k]; }
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j)
                if (i != j && std::fabs(tw.m[i * 4 + j]) > 1e-12) fail("Error while loading \"" + m.file + "\": the grid's to_world must be axis-aligned (scale + translate)");
            double a[3], b[3]; apply_point(tw, lo, a); apply_point(tw, hi, b);
            for (int k = 0; k < 3; ++k) { d.grid_bbox_min[k] = (float) std::min(a[k], b[k]); d.grid_bbox_max[k] = (float) std::max(a[k], b[k]); }
            d.grid_res[0] = g.nx; d.grid_res[1] = g.ny; d.grid_res[2] = g.nz; d.grid_max = g.max;
            out.grids.emplace_back(new std::vector<float>(std::move(g.data)));
            d.grid = out.grids.back()->data();
        } else fail("Error while loading \"" + m.file + "\": medium type \"" + type + "\" is not on the VRL path");
        const int idx = (int) out.scene.media.size();
        out.scene.media.push_back(d); out.media_ids.push_back(id);
        nl_boxes.push_back({ 1e300, 1e300, 1e300, -1e300, -1e300, -1e300 }); nl_box_set.push_back(false);
        if (!id.empty()) medium_index[id] = idx;
        return idx;
    }

    // ---- emitters ----
    int emitter(const Node &e) {
        const std::string id = e.get("id", "");
        if (!id.empty() && emitter_index.count(id)) return emitter_index[id];
        if (e.get("type") != "area") fail("Error while loading \"" + e.file + "\": emitter type \"" + e.get("type") + "\" is not on the VRL path (area emitters and the integrator's laser are)");
        vrlb_emitter d{};
        d.radiance[0] = d.radiance[1] = d.radiance[2] = 1.f;                                                  // area.cpp:32
        prop_rgb(e, "radiance", d.radiance);
        if (const Node *r = prop(e, "radiance")) if (r->get("value", "").find(':') != std::string::npos)
            warn("radiance given as a wavelength table: reduced to RGB by band means (no CIE integration)");
        const int idx = (int) out.scene.emitters.size();
        if (idx >= 255) fail("too many emitters");
        out.scene.emitters.push_back(d);
        if (!id.empty()) emitter_index[id] = idx;
        return idx;
    }

    // ---- shapes ----
    // named indices of refraction: the table of include/mitsuba/render/ior.h:23-47
    double ior_of(const Node &b, const char *name, const char *def) {
        const Node *p = prop(b, name);
        if (p && (p->name == "float" || p->name == "integer")) return to_double(*p, p->get("value"));
        const std::string v = p ? p->get("value") : std::string(def);
        static const std::pair<const char *, double> table[] = {
            { "vacuum", 1.0 }, { "helium", 1.000036 }, { "hydrogen", 1.000132 }, { "air", 1.000277 }, { "carbon dioxide", 1.00045 },
            { "water", 1.3330 }, { "acetone", 1.36 }, { "ethanol", 1.361 }, { "carbon tetrachloride", 1.461 }, { "glycerol", 1.4729 },
            { "benzene", 1.501 }, { "silicone oil", 1.52045 }, { "bromine", 1.661 }, { "water ice", 1.31 }, { "fused quartz", 1.458 },
            { "pyrex", 1.470 }, { "acrylic glass", 1.49 }, { "polypropylene", 1.49 }, { "bk7", 1.5046 }, { "sodium chloride", 1.544 },
            { "amber", 1.55 }, { "pet", 1.5750 }, { "diamond", 2.419 } };
        for (auto &e : table) if (v == e.first) return e.second;
        fail("Error while loading \"" + b.file + "\": unknown index of refraction \"" + v + "\"");
    }
    void add_tri(const double a[3], const double b[3], const double c[3], uint32_t flags, std::vector<float> &v) {
        for (const double *p : { a, b, c }) for (int k = 0; k < 3; ++k) v.push_back((float) p[k]);
        (void) flags;
    }
    void shape(const Node &s) {
        const std::string type = s.get("type");
        Mat4 tw = to_world_of(s);
        std::vector<float> v;
        auto quad = [&](const double p[4][3]) {
            double w[4][3]; for (int i = 0; i < 4; ++i) apply_point(tw, p[i], w[i]);
            add_tri(w[0], w[1], w[2], 0, v); add_tri(w[2], w[3], w[0], 0, v);
        };
        if (type == "obj") {
            float m[16]; for (int i = 0; i < 16; ++i) m[i] = (float) tw.m[i];
            v = read_obj(resolve(base_dir, prop_string(s, "filename", "")), m);
        } else if (type == "rectangle") {                                                                     // rectangle.cpp: [-1,1]^2 in z = 0, normal +z
            const double p[4][3] = { { -1, -1, 0 }, { 1, -1, 0 }, { 1, 1, 0 }, { -1, 1, 0 } };
            quad(p);
        } else if (type == "cube") {                                                                          // cube.cpp: [-1,1]^3, outward normals
            const double c[8][3] = { { -1, -1, -1 }, { 1, -1, -1 }, { 1, 1, -1 }, { -1, 1, -1 }, { -1, -1, 1 }, { 1, -1, 1 }, { 1, 1, 1 }, { -1, 1, 1 } };
            const int f[6][4] = { { 0, 3, 2, 1 }, { 4, 5, 6, 7 }, { 0, 1, 5, 4 }, { 3, 7, 6, 2 }, { 0, 4, 7, 3 }, { 1, 2, 6, 5 } };
            for (auto &q4 : f) { const double p[4][3] = { { c[q4[0]][0], c[q4[0]][1], c[q4[0]][2] }, { c[q4[1]][0], c[q4[1]][1], c[q4[1]][2] },
                                                          { c[q4[2]][0], c[q4[2]][1], c[q4[2]][2] }, { c[q4[3]][0], c[q4[3]][1], c[q4[3]][2] } }; quad(p); }
        } else if (type == "sphere") {                                                                        // sphere.cpp:98-102: to_world * translate(center) * scale(radius)
            double ctr[3] = { 0, 0, 0 };
            if (const Node *c = prop(s, "center")) xyz(*c, 0.0, ctr);
            const double r = prop_float(s, "radius", 1.0);
            Mat4 T = identity(); T.m[0] = T.m[5] = T.m[10] = r; T.m[3] = ctr[0]; T.m[7] = ctr[1]; T.m[11] = ctr[2];
            const Mat4 M = mul(tw, T);
            const int NLAT = 32, NLON = 64;
            auto P = [&](int i, int j, double o[3]) {
                const double th = M_PI * i / NLAT, ph = 2 * M_PI * j / NLON;
                const double p[3] = { std::sin(th) * std::cos(ph), std::sin(th) * std::sin(ph), std::cos(th) };
                apply_point(M, p, o);
            };
            for (int i = 0; i < NLAT; ++i) for (int j = 0; j < NLON; ++j) {
                double a[3], b[3], c[3], d[3]; P(i, j, a); P(i + 1, j, b); P(i + 1, j + 1, c); P(i, j + 1, d);
                if (i != NLAT - 1) add_tri(a, b, c, 0, v);
                if (i != 0) add_tri(a, c, d, 0, v);
            }
            warn("shape type=\"sphere\" is tessellated (" + std::to_string(NLAT) + " x " + std::to_string(NLON) + "); the reference intersects it analytically");
        } else fail("Error while loading \"" + s.file + "\": shape type \"" + type + "\" is not supported by this loader (obj, rectangle, cube, sphere)");

        // surface word: BSDF kind, media on both sides, emitter
        uint32_t flags = VRLB_SURF_BSDF_SMOOTH;
        int interior = -1;
        float refl[3] = { 0.5f, 0.5f, 0.5f };                  // a shape without <bsdf> gets a diffuse one, reflectance 0.5 (shape.cpp, diffuse.cpp:71)
        auto bsdf_kind = [&](const Node &b0) {
            const Node *b = &b0;
            std::string t = b->get("type");
            for (int hop = 0; hop < 4 && (t == "twosided" || t == "bumpmap" || t == "mask"); ++hop) {
                if (t == "twosided") flags |= VRLB_SURF_TWOSIDED;
                const Node *in = child_named(*b, "bsdf", nullptr);
                if (!in) for (const Node &c : b->children) if (c.name == "ref" && deref(c).name == "bsdf") in = &deref(c);
                if (!in) break;
                b = in; t = b->get("type");
            }
            if (t == "null") { flags = (flags & ~VRLB_SURF_BSDF_MASK) | VRLB_SURF_BSDF_NULL; return; }
            if (t == "conductor") {
                // SmoothConductor (conductor.cpp): delta reflection, weight = specular_reflectance x Fresnel.  The descriptor
                // carries one RGB weight: exact for the default eta = 0, k = 1 (Fresnel = 1); a named material or an explicit
                // eta / k is reduced to its normal-incidence value where given as numbers
                flags = (flags & ~VRLB_SURF_BSDF_MASK) | VRLB_SURF_BSDF_MIRROR;
                refl[0] = refl[1] = refl[2] = 1.f;
                prop_rgb(*b, "specular_reflectance", refl);
                float eta[3] = { 0.f, 0.f, 0.f }, k[3] = { 1.f, 1.f, 1.f };
                const bool have = prop_rgb(*b, "eta", eta) | prop_rgb(*b, "k", k);
                if (have) { for (int c = 0; c < 3; ++c) refl[c] *= ((eta[c] - 1) * (eta[c] - 1) + k[c] * k[c]) / ((eta[c] + 1) * (eta[c] + 1) + k[c] * k[c]);
                            warn("bsdf type=\"conductor\": Fresnel term taken at normal incidence"); }
                if (prop(*b, "material") && prop_string(*b, "material", "none") != "none")
                    warn("bsdf type=\"conductor\": material \"" + prop_string(*b, "material", "") + "\" has no spectral table here; Fresnel term taken as 1");
                return;
            }
            if (t == "dielectric") {
                // SmoothDielectric (dielectric.cpp:  int_ior default bk7, ext_ior default air)
                flags = (flags & ~VRLB_SURF_BSDF_MASK) | VRLB_SURF_BSDF_DIELECTRIC;
                float sr[3] = { 1.f, 1.f, 1.f }, st[3] = { 1.f, 1.f, 1.f };
                prop_rgb(*b, "specular_reflectance", sr); prop_rgb(*b, "specular_transmittance", st);
                if (sr[0] != sr[1] || sr[0] != sr[2] || st[0] != st[1] || st[0] != st[2]) warn("bsdf type=\"dielectric\": coloured specular_reflectance / transmittance reduced to the red channel");
                refl[0] = (float) ior_of(*b, "int_ior", "bk7") / (float) ior_of(*b, "ext_ior", "air");   /* m_eta = int_ior / ext_ior in ScalarFloat, dielectric.cpp */ refl[1] = sr[0]; refl[2] = st[0];
                return;
            }
            if (t == "thindielectric") {
                warn("bsdf type=\"thindielectric\" is not on the VRL path; the surface loads as an occluder");
                refl[0] = refl[1] = refl[2] = 0.f;              // light-tracing particles end there
                return;
            }
            // what the light tracer needs of a smooth BSDF: the weight of SmoothDiffuse::sample = its reflectance
            if (t != "diffuse") warn("bsdf type=\"" + t + "\": light-tracing particles are reflected with a diffuse lobe of its reflectance");
            refl[0] = refl[1] = refl[2] = 0.5f;
            const Node *r = prop(*b, "reflectance");
            if (!r) r = prop(*b, "diffuse_reflectance");
            if (!r) return;
            const std::string v = r->get("value", "");
            if ((r->name == "rgb" || r->name == "spectrum" || r->name == "float" || r->name == "color") && v.find(':') == std::string::npos && !v.empty()) prop_rgb(*b, r->get("name").c_str(), refl);
            else if (r->name == "spectrum" && v.find(':') != std::string::npos) {
                prop_rgb(*b, r->get("name").c_str(), refl);
                warn("reflectance given as a wavelength table: reduced to RGB by band means (no CIE integration)");
            } else warn("bsdf \"" + b->get("id", t.c_str()) + "\": textured reflectance; light-tracing particles see the default 0.5");
        };
        for (const Node &c : s.children) {
            if (c.name == "bsdf") bsdf_kind(c);
            else if (c.name == "medium") { int m = medium(c); const std::string nm = c.get("name", "interior");
                                           if (nm == "interior") { flags |= VRLB_SURF_INTERIOR(m); interior = m; } else flags |= VRLB_SURF_EXTERIOR(m); }
            else if (c.name == "emitter") { flags |= VRLB_SURF_EMITTER | VRLB_SURF_EMITTER_ID(emitter(c)); }
            else if (c.name == "ref") {
                const Node &t = deref(c);
                if (t.name == "bsdf") bsdf_kind(t);
                else if (t.name == "medium") { int m = medium(t); const std::string nm = c.get("name", "interior");
                                               if (nm == "interior") { flags |= VRLB_SURF_INTERIOR(m); interior = m; } else flags |= VRLB_SURF_EXTERIOR(m); }
                else if (t.name == "emitter") flags |= VRLB_SURF_EMITTER | VRLB_SURF_EMITTER_ID(emitter(t));
                else fail("Error while loading \"" + s.file + "\": a shape cannot reference a <" + t.name + ">");
            }
        }
        // Emitter::medium() is the exterior medium of the shape the emitter sits on (shape.cpp: the area emitter inherits
        // it); particles start there (vrl.cpp:152)
        if ((flags & VRLB_SURF_EMITTER) && VRLB_SURF_GET_EXTERIOR(flags) >= 0) out.scene.emission_medium = VRLB_SURF_GET_EXTERIOR(flags);
        const bool reflects = (flags & VRLB_SURF_BSDF_MASK) != VRLB_SURF_BSDF_NULL && !((flags & VRLB_SURF_EMITTER) && (flags & VRLB_SURF_BSDF_MASK) == VRLB_SURF_BSDF_SMOOTH);
        for (size_t i = 0; i < v.size() / 9; ++i) {
            out.scene.tri_flags.push_back(flags);
            for (int k = 0; k < 3; ++k) out.scene.tri_reflectance.push_back(reflects ? refl[k] : 0.f);
        }
        out.scene.tri_verts.insert(out.scene.tri_verts.end(), v.begin(), v.end());
        if (interior >= 0) {                                   // NonLinearMedia::build(bbox of the shape), nonlinear.cpp:62-89
            auto &bx = nl_boxes[(size_t) interior]; nl_box_set[(size_t) interior] = true;
            for (size_t i = 0; i < v.size(); ++i) { int k = (int) (i % 3); bx[(size_t) k] = std::min(bx[(size_t) k], (double) v[i]); bx[(size_t) (3 + k)] = std::max(bx[(size_t) (3 + k)], (double) v[i]); }
        }
    }

    // ---- sensor: src/sensors/perspective.cpp:40-80, sensor.h:195-231 (perspective_projection), sensor.cpp (parse_fov) ----
    void sensor(const Node &s) {
        if (s.get("type") != "perspective") fail("Error while loading \"" + s.file + "\": sensor type \"" + s.get("type") + "\" is not on the VRL path (perspective is)");
        const Node *film = child_named(s, "film", nullptr), *sampler = child_named(s, "sampler", nullptr);
        const int w = film ? (int) prop_int(*film, "width", 768) : 768, h = film ? (int) prop_int(*film, "height", 576) : 576;   // film.cpp:15-16
        { const Node *rf = film ? child_named(*film, "rfilter", nullptr) : nullptr;                          // film.cpp:36-48: no <rfilter> = gaussian
          const std::string rt = rf ? rf->get("type") : std::string("gaussian");
          if (rt != "box") warn("rfilter type=\"" + rt + "\"" + (rf ? "" : " (the film's default)") + ": the device film splats with the box filter (imageblock.cpp:162-169 fast path)"); }
        if (film) { const std::string cf = prop_string(*film, "component_format", "float16");               // hdrfilm.cpp:108-160
                    out.film_half = cf == "float16";
                    if (cf != "float16" && cf != "float32") warn("film component_format=\"" + cf + "\": written as float32"); }
        out.sensor.spp = sampler ? (int32_t) prop_int(*sampler, "sample_count", 4) : 4;                      // sampler.cpp:11
        out.sensor.sampler_seed = sampler ? (uint64_t) prop_int(*sampler, "seed", 0) : 0;
        if (sampler && sampler->get("type") != "independent") warn("sampler type=\"" + sampler->get("type") + "\": the path addresses independent PCG32 streams");
        const double aspect = (double) w / (double) h, near = prop_float(s, "near_clip", 1e-2), far = prop_float(s, "far_clip", 1e4);
        double fov = prop_float(s, "fov", -1.0);
        if (fov < 0) fail("Error while loading \"" + s.file + "\": the perspective sensor needs \"fov\" (focal_length is not supported by this loader)");
        std::string axis = prop_string(s, "fov_axis", "x");
        if (axis == "smaller") axis = aspect > 1 ? "y" : "x"; else if (axis == "larger") axis = aspect > 1 ? "x" : "y";
        double xfov = fov;
        if (axis == "y") xfov = 2 * std::atan(std::tan(0.5 * fov * M_PI / 180) * aspect) * 180 / M_PI;
        else if (axis == "diagonal") { double diag = 2 * std::tan(0.5 * fov * M_PI / 180), width = diag / std::sqrt(1 + 1 / (aspect * aspect)); xfov = 2 * std::atan(width * 0.5) * 180 / M_PI; }
        else if (axis != "x") fail("Error while loading \"" + s.file + "\": fov_axis must be x, y, diagonal, smaller or larger");
        // camera_to_sample = scale(-0.5, -0.5 aspect, 1) * translate(-1, -1/aspect, 0) * perspective(xfov, near, far); we ship its inverse
        const double cot = 1.0 / std::tan(0.5 * xfov * M_PI / 180), recip = 1.0 / (far - near);
        Mat4 P{}; P.m[0] = cot; P.m[5] = cot; P.m[10] = far * recip; P.m[11] = -near * far * recip; P.m[14] = 1.0;
        Mat4 T = identity(); T.m[3] = -1.0; T.m[7] = -1.0 / aspect;
        Mat4 S = identity(); S.m[0] = -0.5; S.m[5] = -0.5 * aspect;
        const Mat4 c2s = mul(S, mul(T, P));
        // inverse of [[a 0 0 b],[0 c 0 d],[0 0 e f],[0 0 1 0]]: x = (X - b z)/a ..., solved in closed form
        const double a = c2s.m[0], b = c2s.m[2], c = c2s.m[5], d = c2s.m[6], e = c2s.m[10], f = c2s.m[11];
        // rows of c2s: (a 0 b 0) (0 c d 0) (0 0 e f) (0 0 1 0)  [the translate lands in the z column after the projective row]
        Mat4 I{};
        I.m[0] = 1 / a; I.m[3] = -b / a;          // x = (X - b W) / a        with W = z
        I.m[5] = 1 / c; I.m[7] = -d / c;          // y = (Y - d W) / c
        I.m[11] = 1.0;                            // z = W
        I.m[14] = 1 / f; I.m[15] = -e / f;        // w = (Z - e W) / f
        vrlb_camera &cam = out.sensor.camera;
        for (int i = 0; i < 16; ++i) cam.sample_to_camera[i] = (float) I.m[i];
        const Mat4 tw = to_world_of(s);
        for (int i = 0; i < 16; ++i) cam.to_world[i] = (float) tw.m[i];
        cam.near_clip = (float) near; cam.far_clip = (float) far; cam.width = w; cam.height = h; cam.medium = -1;
        for (const Node &ch : s.children) {
            if (ch.name == "medium") cam.medium = medium(ch);
            else if (ch.name == "ref" && deref(ch).name == "medium") cam.medium = medium(deref(ch));
        }
    }

    void integrator(const Node &n) {
        out.integrator_type = n.get("type");
        out.integrator = Properties(out.integrator_type);
        for (const Node &c : n.children) {
            const std::string k = c.get("name", "");
            if (k.empty()) fail("Error while loading \"" + n.file + "\": <" + c.name + "> inside <integrator> needs a name");
            if (c.name == "integer") out.integrator.set_int(k, to_int(c, c.get("value")));
            else if (c.name == "float") out.integrator.set_float(k, to_double(c, c.get("value")));
            else if (c.name == "boolean") { std::string v = c.get("value"); if (v != "true" && v != "false") fail("Error while loading \"" + n.file + "\": could not parse boolean value \"" + v + "\"");
                                            out.integrator.set_bool(k, v == "true"); }
            else if (c.name == "string") out.integrator.set_string(k, c.get("value"));
            else if (c.name == "vector" || c.name == "point") { double v[3]; xyz(c, 0.0, v); out.integrator.set_vector3f(k, Vec3{ (float) v[0], (float) v[1], (float) v[2] }); }
            else fail("Error while loading \"" + n.file + "\": unsupported property tag <" + c.name + "> in <integrator>");
        }
    }
};

}  // namespace

LoadedScene load_scene_xml(const std::string &path, const std::map<std::string, std::string> &params_in) {
    std::string text = read_text(path);
    Node root = Parser(text, path).document();
    if (root.name != "scene") fail("Error while loading \"" + path + "\": root element must be <scene>");
    std::vector<std::string> notes;
    upgrade_tree(root, path, notes);              // scene versions before 2.0 (xml.cpp:350-392); checks the version attribute
    Params params;
    for (auto &p : params_in) set_param(params, p.first, p.second, false);
    expand(root, params, dir_of(path), 0, notes);

    Builder b;
    for (auto &w : notes) b.warn(w);
    b.base_dir = dir_of(path);
    b.index_ids(root);
    // media first (ids are referenced by shapes and the sensor), in document order so that ids are stable
    for (const Node &c : root.children) if (c.name == "medium") b.medium(c);
    for (const Node &c : root.children) if (c.name == "emitter" && c.get("type") == "area") b.emitter(c);
    bool have_sensor = false;
    for (const Node &c : root.children) {
        if (c.name == "shape") b.shape(c);
        else if (c.name == "sensor") { if (have_sensor) b.warn("more than one sensor: the first is used"); else { b.sensor(c); have_sensor = true; } }
        else if (c.name == "integrator") b.integrator(c);
        else if (c.name == "emitter" && c.get("type") != "area")
            b.warn("emitter type=\"" + c.get("type") + "\" is ignored: VRLs are traced from area emitters or the integrator's laser (vrl.cpp:139-162)");
        else if (c.name == "bsdf" || c.name == "medium" || c.name == "emitter" || c.name == "texture" || c.name == "phase") {}
        else fail("Error while loading \"" + path + "\": unsupported top-level tag <" + c.name + ">");
    }
    if (!have_sensor) fail("Error while loading \"" + path + "\": the scene has no sensor");
    for (size_t m = 0; m < b.out.scene.media.size(); ++m) {
        vrlb_medium &d = b.out.scene.media[m];
        if (d.kind != VRLB_MEDIUM_NONLINEAR) continue;
        if (!b.nl_box_set[m]) fail("Error while loading \"" + path + "\": nonlinear medium \"" + b.out.media_ids[m] + "\" is not the interior of any shape (its IOR grid has no bounds)");
        for (int k = 0; k < 3; ++k) { d.nl_bbox_min[k] = (float) b.nl_boxes[m][(size_t) k]; d.nl_bbox_max[k] = (float) b.nl_boxes[m][(size_t) (3 + k)]; }
    }
    return std::move(b.out);
}

}  // namespace vrlhost
