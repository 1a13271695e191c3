// vrl_xml.cpp -- see vrl_xml.h.  Part 1: a small XML reader, parameter substitution, includes, transforms.
#include "vrl_xml.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>

#include "vrl_io.h"

namespace vrlhost {
namespace {

[[noreturn]] void fail(const std::string &msg) { throw std::runtime_error(msg); }

struct Node {
    std::string name;
    std::vector<std::pair<std::string, std::string>> attrs;
    std::vector<Node> children;
    std::string file;
    const std::string *attr(const char *k) const {
        for (auto &a : attrs) if (a.first == k) return &a.second;
        return nullptr;
    }
    std::string get(const char *k, const char *def = nullptr) const {
        if (auto *v = attr(k)) return *v;
        if (def) return def;
        fail("Error while loading \"" + file + "\": missing attribute \"" + k + "\" in <" + name + ">");
    }
};

// ---- XML text -> Node tree (elements and attributes only; text content, comments, declarations are skipped) ----
struct Parser {
    const std::string &s; size_t i = 0; std::string file;
    Parser(const std::string &text, std::string f) : s(text), file(std::move(f)) {}
    [[noreturn]] void err(const std::string &m) const {
        size_t line = 1 + (size_t) std::count(s.begin(), s.begin() + (long) std::min(i, s.size()), '\n');
        fail("Error while loading \"" + file + "\" (line " + std::to_string(line) + "): " + m);
    }
    void skip_ws() { while (i < s.size() && std::isspace((unsigned char) s[i])) ++i; }
    bool starts(const char *t) const { return s.compare(i, std::strlen(t), t) == 0; }
    void skip_misc() {        // whitespace, text, comments, <? ?>, <! >
        for (;;) {
            while (i < s.size() && s[i] != '<') ++i;
            if (i >= s.size()) return;
            if (starts("<!--")) { size_t e = s.find("-->", i + 4); if (e == std::string::npos) err("unterminated comment"); i = e + 3; }
            else if (starts("<?")) { size_t e = s.find("?>", i + 2); if (e == std::string::npos) err("unterminated declaration"); i = e + 2; }
            else if (starts("<!")) { size_t e = s.find('>', i + 2); if (e == std::string::npos) err("unterminated declaration"); i = e + 1; }
            else return;
        }
    }
    std::string ident() {
        size_t b = i;
        while (i < s.size() && (std::isalnum((unsigned char) s[i]) || s[i] == '_' || s[i] == '-' || s[i] == ':' || s[i] == '.')) ++i;
        if (b == i) err("expected a name");
        return s.substr(b, i - b);
    }
    static std::string unescape(std::string v) {
        const std::pair<const char *, const char *> ent[] = { { "&lt;", "<" }, { "&gt;", ">" }, { "&quot;", "\"" }, { "&apos;", "'" }, { "&amp;", "&" } };
        for (auto &e : ent) for (size_t p = 0; (p = v.find(e.first, p)) != std::string::npos; p += 1) v.replace(p, std::strlen(e.first), e.second);
        return v;
    }
    Node element() {          // at '<' of an opening tag
        Node n; n.file = file;
        ++i; n.name = ident();
        for (;;) {
            skip_ws();
            if (i >= s.size()) err("unexpected end of file in <" + n.name + ">");
            if (s[i] == '/') { if (!starts("/>")) err("malformed tag"); i += 2; return n; }
            if (s[i] == '>') { ++i; break; }
            std::string k = ident();
            skip_ws(); if (i >= s.size() || s[i] != '=') err("expected '=' after attribute \"" + k + "\"");
            ++i; skip_ws();
            if (i >= s.size() || (s[i] != '"' && s[i] != '\'')) err("expected a quoted value for \"" + k + "\"");
            char q = s[i++]; size_t e = s.find(q, i);
            if (e == std::string::npos) err("unterminated attribute value");
            n.attrs.emplace_back(k, unescape(s.substr(i, e - i)));
            i = e + 1;
        }
        for (;;) {
            skip_misc();
            if (i >= s.size()) err("missing </" + n.name + ">");
            if (starts("</")) {
                i += 2; std::string c = ident(); skip_ws();
                if (c != n.name) err("mismatched closing tag </" + c + "> for <" + n.name + ">");
                if (i >= s.size() || s[i] != '>') err("malformed closing tag");
                ++i; return n;
            }
            n.children.push_back(element());
        }
    }
    Node document() {
        skip_misc();
        if (i >= s.size()) err("no root element");
        Node r = element();
        return r;
    }
};

std::string read_text(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) fail("Error while loading \"" + path + "\": file not found");
    std::ostringstream o; o << f.rdbuf();
    return o.str();
}
std::string dir_of(const std::string &path) {
    size_t p = path.find_last_of("/\\");
    return p == std::string::npos ? std::string(".") : path.substr(0, p);
}
std::string resolve(const std::string &base_dir, const std::string &name) {
    if (!name.empty() && (name[0] == '/' || (name.size() > 1 && name[1] == ':'))) return name;
    return base_dir + "/" + name;
}

// ---- $param substitution (xml.cpp:618-649: longest names first, an unresolved '$' is an error) ----
using Params = std::vector<std::pair<std::string, std::string>>;
void substitute(Node &n, const Params &params) {
    for (auto &a : n.attrs) {
        if (a.second.find('$') == std::string::npos) continue;
        for (auto &p : params) {
            const std::string key = "$" + p.first;
            for (size_t pos = 0; (pos = a.second.find(key, pos)) != std::string::npos; pos += p.second.size()) a.second.replace(pos, key.size(), p.second);
        }
        if (a.second.find('$') != std::string::npos)
            fail("Error while loading \"" + n.file + "\": undefined parameter in <" + n.name + " " + a.first + "=\"" + a.second + "\">");
    }
}
void set_param(Params &params, const std::string &k, const std::string &v, bool keep_existing) {
    for (auto &p : params) if (p.first == k) { if (!keep_existing) p.second = v; return; }
    params.emplace_back(k, v);
    std::stable_sort(params.begin(), params.end(), [](auto &a, auto &b) { return a.first.size() > b.first.size(); });
}

bool upgrade_tree(Node &root, const std::string &path, std::vector<std::string> &notes);

// Depth-first: <default> defines a parameter unless the caller already set it, <include> splices the children of the
// included file's root in place (xml.cpp:651-690), every other tag gets its attributes substituted.
void expand(Node &n, Params &params, const std::string &base_dir, int depth, std::vector<std::string> &notes) {
    if (depth > 16) fail("Error while loading \"" + n.file + "\": includes nested too deeply");
    std::vector<Node> out;
    for (Node &c : n.children) {
        substitute(c, params);
        if (c.name == "default") { set_param(params, c.get("name"), c.get("value"), true); continue; }
        if (c.name == "include") {
            const std::string path = resolve(base_dir, c.get("filename"));
            std::string text = read_text(path);
            Node root = Parser(text, path).document();
            if (root.name != "scene") fail("Error while loading \"" + path + "\": root element of an included file must be <scene>");
            upgrade_tree(root, path, notes);
            expand(root, params, dir_of(path), depth + 1, notes);
            for (Node &g : root.children) out.push_back(std::move(g));
            continue;
        }
        expand(c, params, base_dir, depth, notes);
        out.push_back(std::move(c));
    }
    n.children = std::move(out);
}

// ---- scene versions before 2.0 (xml.cpp:350-392 upgrade_tree, run per file like the reference does) ----
// The reference's own in-memory upgrade: every name="..." except on <default> goes from camelCase to underscore_case,
// <lookAt> becomes <lookat>, ids with a leading underscore are renamed, diffuse_reflectance of a diffuse BSDF becomes
// reflectance.  On top of that, what the 0.x scenes of data/FBD15ALVRL_scenes need and no plugin of the fork provides
// ("reusable after schema translation"), each reported in `notes`:
//   medium sigma_s + sigma_a            -> sigma_t = sigma_s + sigma_a, albedo = sigma_s / sigma_t
//   shape with a medium and no BSDF     -> index-matched boundary (0.x Shape::configure), i.e. <bsdf type="null"/>
//   gridvolume without a to_world       -> use_grid_bbox = true (the 0.x plugin places the grid by the file's bounding box)
//   <integrator type="vrl"> properties  -> those of the LightSlice implementation these scenes were written for are dropped;
//                                          VRLIntegrator then runs on its defaults (or the caller's overrides)
int version_major(const Node &root, const std::string &path) {
    const std::string v = root.get("version", "");
    if (v.empty()) fail("Error while loading \"" + path + "\": missing version attribute in root element <scene>");
    size_t used = 0; int major = -1;
    try { major = std::stoi(v, &used); } catch (...) {}
    if (major < 0 || (used < v.size() && v[used] != '.')) fail("Error while loading \"" + path + "\": could not parse scene version \"" + v + "\"");
    return major;
}
std::string underscore_case(std::string name) {         // the loop of xml.cpp:365-375, including its skip after a hump
    for (size_t i = 0; i + 1 < name.length(); ++i) {
        if (std::islower((unsigned char) name[i]) && std::isupper((unsigned char) name[i + 1])) {
            name = name.substr(0, i + 1) + "_" + name.substr(i + 1);
            i += 2;
            while (i < name.length() && std::isupper((unsigned char) name[i])) { name[i] = (char) std::tolower((unsigned char) name[i]); ++i; }
        }
    }
    return name;
}
const char *const kVrlProps[] = {       // what VRLIntegrator's constructor queries (vrl_integrator.cpp <- vrl.cpp:38-104)
    "block_size", "caustic_lookup_radius_relative", "caustic_lookup_size", "caustic_photons", "device", "dice_vrl",
    "global_lookup_radius_relative", "global_lookup_size", "global_photons", "hide_emitters", "laser_direction", "laser_origin",
    "lightcut_samples", "long_vrl", "math_mode", "max_depth", "min_vrl_length", "rr_depth", "rr_vrl", "samples_per_pass",
    "samples_per_query", "scale_rr", "shoot_center", "stochastic_lightcut", "target_vrls", "threshold_better_dist", "threshold_error",
    "timeout", "use_direct_illum", "use_laser", "use_light_cut", "use_nl_atomic_query", "use_non_linear", "use_non_linear_camera",
    "use_uniform_sampling", "volume_lookup_radius_relative", "volume_photons" };
void set_attr(Node &n, const char *k, const std::string &v) { for (auto &a : n.attrs) if (a.first == k) { a.second = v; return; } n.attrs.emplace_back(k, v); }
void upgrade_names(Node &n, std::vector<std::string> &notes) {
    if (n.name == "lookAt") n.name = "lookat";
    if (n.name != "default") if (auto *nm = n.attr("name")) if (!nm->empty()) set_attr(n, "name", underscore_case(*nm));
    if (auto *id = n.attr("id")) if (!id->empty() && (*id)[0] == '_') {
        const std::string renamed = "ID" + *id + "__UPGR";
        notes.push_back("changing identifier: \"" + *id + "\" -> \"" + renamed + "\"");
        set_attr(n, "id", renamed);
    }
    if (n.name == "bsdf" && n.get("type", "") == "diffuse")
        for (Node &c : n.children) if (c.get("name", "") == "diffuse_reflectance") set_attr(c, "name", "reflectance");
    for (Node &c : n.children) upgrade_names(c, notes);
}
void translate_legacy(Node &n, std::vector<std::string> &notes) {
    auto find = [&](const char *nm) -> Node * { for (Node &c : n.children) if (c.name != "ref" && c.get("name", "") == nm) return &c; return nullptr; };
    if (n.name == "medium" && find("sigma_s") && find("sigma_a") && !find("sigma_t")) {
        Node *ns = find("sigma_s"), *na = find("sigma_a");
        auto comps = [&](const Node &c) {
            const std::string v = c.get("value");
            if (v.find(':') != std::string::npos) fail("Error while loading \"" + n.file + "\": medium coefficient \"" + c.get("name") + "\" is a wavelength table; only constants are translated");
            std::vector<double> r; std::string tok;
            for (char ch : v + ",") { if (ch == ',' || std::isspace((unsigned char) ch)) { if (!tok.empty()) { r.push_back(std::stod(tok)); tok.clear(); } } else tok += ch; }
            if (r.size() == 1) r = { r[0], r[0], r[0] };
            if (r.size() != 3) fail("Error while loading \"" + n.file + "\": medium coefficient \"" + c.get("name") + "\" needs 1 or 3 components");
            return r;
        };
        const auto s = comps(*ns), a = comps(*na);
        std::ostringstream st, al, note; st.precision(17); al.precision(17); note.precision(6);       // 17 digits: the double survives the round trip
        for (int k = 0; k < 3; ++k) {
            const double t = s[(size_t) k] + a[(size_t) k], alb = t > 0 ? s[(size_t) k] / t : 0.0;
            st << (k ? ", " : "") << t; al << (k ? ", " : "") << alb;
            note << (k ? ", " : "") << t << " / " << alb;
        }
        Node t; t.name = "rgb"; t.file = n.file; t.attrs = { { "name", "sigma_t" }, { "value", st.str() } };
        Node b; b.name = "rgb"; b.file = n.file; b.attrs = { { "name", "albedo" }, { "value", al.str() } };
        std::vector<Node> kept;
        for (Node &c : n.children) if (&c != ns && &c != na) kept.push_back(std::move(c));
        kept.push_back(std::move(t)); kept.push_back(std::move(b));
        n.children = std::move(kept);
        notes.push_back("medium \"" + n.get("id", "") + "\": sigma_s / sigma_a translated to sigma_t / albedo = " + note.str());
    }
    if (n.name == "shape") {
        bool has_bsdf = false, has_medium = false, has_emitter = false;
        for (const Node &c : n.children) {
            if (c.name == "bsdf") has_bsdf = true;
            else if (c.name == "medium") has_medium = true;
            else if (c.name == "emitter") has_emitter = true;
            else if (c.name == "ref") { const std::string nm = c.get("name", ""); if (nm == "interior" || nm == "exterior") has_medium = true; else has_bsdf = true; }
        }
        if (has_medium && !has_bsdf && !has_emitter) {
            Node b; b.name = "bsdf"; b.file = n.file; b.attrs = { { "type", "null" } };
            n.children.push_back(std::move(b));
            notes.push_back("0.x shape with a medium and no BSDF: loaded as an index-matched boundary (null BSDF), as 0.x Shape::configure did");
        }
    }
    if (n.name == "volume" && n.get("type", "") == "gridvolume" && !find("use_grid_bbox") && !find("to_world")) {
        // a 0.x gridvolume sits where its file's bounding box says (0.x gridvolume.cpp); Mitsuba 2 asks for use_grid_bbox
        Node b; b.name = "boolean"; b.file = n.file; b.attrs = { { "name", "use_grid_bbox" }, { "value", "true" } };
        n.children.push_back(std::move(b));
        notes.push_back("0.x gridvolume: placed by the bounding box stored in its file (use_grid_bbox = true)");
    }
    if (n.name == "integrator" && n.get("type", "") == "vrl") {
        std::vector<Node> kept;
        std::string dropped;
        for (Node &c : n.children) {
            const std::string nm = c.get("name", "");
            bool known = false;
            for (const char *k : kVrlProps) known = known || nm == k;
            if (known) kept.push_back(std::move(c));
            else dropped += (dropped.empty() ? "" : ", ") + nm;
        }
        n.children = std::move(kept);
        if (!dropped.empty()) notes.push_back("0.x <integrator type=\"vrl\"> properties of another VRL implementation are dropped: " + dropped);
    }
    for (Node &c : n.children) translate_legacy(c, notes);
}
// returns true when the file was upgraded
bool upgrade_tree(Node &root, const std::string &path, std::vector<std::string> &notes) {
    const int major = version_major(root, path);
    if (major >= 2) return false;
    notes.push_back("\"" + path + "\": in-memory version upgrade (v" + root.get("version") + " -> v2.0.0)");
    upgrade_names(root, notes);
    translate_legacy(root, notes);
    return true;
}

// ---- numbers, vectors, transforms ----
double to_double(const Node &n, const std::string &v) {
    try {
        size_t used = 0;
        double d = std::stod(v, &used);
        while (used < v.size() && std::isspace((unsigned char) v[used])) ++used;
        if (used == v.size()) return d;
    } catch (...) {}
    fail("Error while loading \"" + n.file + "\": could not parse floating point value \"" + v + "\" in <" + n.name + ">");
}
int64_t to_int(const Node &n, const std::string &v) {
    try {
        size_t used = 0;
        long long d = std::stoll(v, &used);
        while (used < v.size() && std::isspace((unsigned char) v[used])) ++used;
        if (used == v.size()) return d;
    } catch (...) {}
    fail("Error while loading \"" + n.file + "\": could not parse integer value \"" + v + "\" in <" + n.name + ">");
}
std::vector<double> to_list(const Node &n, const std::string &v) {
    std::vector<double> r; std::string tok;
    for (char ch : v + ",") {
        if (ch == ',' || std::isspace((unsigned char) ch)) { if (!tok.empty()) { r.push_back(to_double(n, tok)); tok.clear(); } }
        else tok += ch;
    }
    return r;
}
// x="" y="" z="" or value="a, b, c" or value="a" (all three), as parse_vector / parse_named_vector in xml.cpp
void xyz(const Node &n, double def, double out[3]) {
    if (auto *v = n.attr("value")) {
        auto l = to_list(n, *v);
        if (l.size() == 1) { out[0] = out[1] = out[2] = l[0]; return; }
        if (l.size() == 3) { out[0] = l[0]; out[1] = l[1]; out[2] = l[2]; return; }
        fail("Error while loading \"" + n.file + "\": <" + n.name + " value=\"" + *v + "\"> needs 1 or 3 components");
    }
    const char *k[3] = { "x", "y", "z" };
    for (int i = 0; i < 3; ++i) out[i] = n.attr(k[i]) ? to_double(n, *n.attr(k[i])) : def;
}

struct Mat4 { double m[16]; };
Mat4 identity() { Mat4 r{}; for (int i = 0; i < 4; ++i) r.m[i * 5] = 1.0; return r; }
Mat4 mul(const Mat4 &a, const Mat4 &b) {
    Mat4 r{};
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += a.m[i * 4 + k] * b.m[k * 4 + j]; r.m[i * 4 + j] = s; }
    return r;
}
void apply_point(const Mat4 &t, const double p[3], double out[3]) {
    double w = t.m[12] * p[0] + t.m[13] * p[1] + t.m[14] * p[2] + t.m[15];
    for (int i = 0; i < 3; ++i) out[i] = (t.m[i * 4] * p[0] + t.m[i * 4 + 1] * p[1] + t.m[i * 4 + 2] * p[2] + t.m[i * 4 + 3]) / w;
}
void norm3(double v[3]) { double l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); for (int i = 0; i < 3; ++i) v[i] /= l; }
void cross3(const double a[3], const double b[3], double o[3]) { o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0]; }

// <transform>: operations apply in document order, i.e. each one multiplies from the left (xml.cpp:806-905)
Mat4 parse_transform(const Node &t) {
    Mat4 M = identity();
    for (const Node &op : t.children) {
        Mat4 T = identity();
        if (op.name == "translate") { double v[3]; xyz(op, 0.0, v); T.m[3] = v[0]; T.m[7] = v[1]; T.m[11] = v[2]; }
        else if (op.name == "scale") { double v[3]; xyz(op, 1.0, v); T.m[0] = v[0]; T.m[5] = v[1]; T.m[10] = v[2]; }
        else if (op.name == "rotate") {
            double a[3]; xyz(op, 0.0, a);
            if (a[0] == 0 && a[1] == 0 && a[2] == 0) fail("Error while loading \"" + op.file + "\": <rotate> needs an axis");
            norm3(a);
            const double ang = to_double(op, op.get("angle")) * M_PI / 180.0, c = std::cos(ang), s = std::sin(ang), k = 1 - c;
            const double R[9] = { c + a[0] * a[0] * k, a[0] * a[1] * k - a[2] * s, a[0] * a[2] * k + a[1] * s,
                                  a[1] * a[0] * k + a[2] * s, c + a[1] * a[1] * k, a[1] * a[2] * k - a[0] * s,
                                  a[2] * a[0] * k - a[1] * s, a[2] * a[1] * k + a[0] * s, c + a[2] * a[2] * k };
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) T.m[i * 4 + j] = R[i * 3 + j];
        } else if (op.name == "lookat") {
            auto o = to_list(op, op.get("origin")), g = to_list(op, op.get("target")), u = to_list(op, op.get("up", "0, 1, 0"));
            if (o.size() != 3 || g.size() != 3 || u.size() != 3) fail("Error while loading \"" + op.file + "\": <lookat> needs 3-vectors");
            double d[3] = { g[0] - o[0], g[1] - o[1], g[2] - o[2] }, left[3], nu[3], up[3] = { u[0], u[1], u[2] };
            norm3(d); cross3(up, d, left); norm3(left); cross3(d, left, nu);               // Transform::look_at, transform.h
            for (int i = 0; i < 3; ++i) { T.m[i * 4] = left[i]; T.m[i * 4 + 1] = nu[i]; T.m[i * 4 + 2] = d[i]; T.m[i * 4 + 3] = o[(size_t) i]; }
        } else if (op.name == "matrix") {
            auto l = to_list(op, op.get("value"));
            if (l.size() == 16) for (int i = 0; i < 16; ++i) T.m[i] = l[(size_t) i];
            else if (l.size() == 9) for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) T.m[i * 4 + j] = l[(size_t) (i * 3 + j)];
            else fail("Error while loading \"" + op.file + "\": <matrix> needs 9 or 16 values");
        } else fail("Error while loading \"" + op.file + "\": unsupported transform operation <" + op.name + ">");
        M = mul(T, M);
    }
    return M;
}
const Node *child_named(const Node &n, const char *tag, const char *name_attr) {
    for (const Node &c : n.children) if (c.name == tag && (!name_attr || c.get("name", "") == name_attr)) return &c;
    return nullptr;
}
Mat4 to_world_of(const Node &n) {
    const Node *t = child_named(n, "transform", "to_world");
    return t ? parse_transform(*t) : identity();
}
// typed property lookup inside a plugin tag: <float name="fov" value=".."/> etc.
const Node *prop(const Node &n, const char *name) {
    for (const Node &c : n.children) if (c.attr("name") && *c.attr("name") == name && c.name != "ref") return &c;
    return nullptr;
}
double prop_float(const Node &n, const char *name, double def) {
    const Node *p = prop(n, name);
    if (!p) return def;
    if (p->name != "float" && p->name != "integer") fail("Error while loading \"" + n.file + "\": property \"" + name + "\" of <" + n.name + "> has the wrong type");
    return to_double(*p, p->get("value"));
}
int64_t prop_int(const Node &n, const char *name, int64_t def) {
    const Node *p = prop(n, name);
    if (!p) return def;
    if (p->name != "integer") fail("Error while loading \"" + n.file + "\": property \"" + name + "\" of <" + n.name + "> has the wrong type");
    return to_int(*p, p->get("value"));
}
std::string prop_string(const Node &n, const char *name, const std::string &def) {
    const Node *p = prop(n, name);
    return p ? p->get("value") : def;
}
bool prop_bool(const Node &n, const char *name, bool def) {
    const Node *p = prop(n, name);
    if (!p) return def;
    std::string v = p->get("value");
    if (v == "true") return true;
    if (v == "false") return false;
    fail("Error while loading \"" + n.file + "\": could not parse boolean value \"" + v + "\"");
}
// constant colour: <rgb>/<spectrum>/<float name=..> with 1 or 3 components
bool prop_rgb(const Node &n, const char *name, float out[3]) {
    const Node *p = prop(n, name);
    if (!p) return false;
    if (p->name != "rgb" && p->name != "spectrum" && p->name != "float" && p->name != "color") return false;
    if (p->get("value").find(':') != std::string::npos) {
        // wavelength:value table (the 0.x scenes and data/cbox): mean over the red / green / blue thirds of the visible range
        // (no CIE integration; callers report it in `warnings`)
        double sum[3] = { 0, 0, 0 }; int cnt[3] = { 0, 0, 0 }; std::string tok;
        for (char ch : p->get("value") + ",") {
            if (ch != ',') { tok += ch; continue; }
            const size_t c = tok.find(':');
            if (c != std::string::npos) { const double wl = to_double(*p, tok.substr(0, c)), val = to_double(*p, tok.substr(c + 1));
                                          const int band = wl < 490 ? 2 : wl < 580 ? 1 : 0; sum[band] += val; ++cnt[band]; }
            tok.clear();
        }
        const double all = (sum[0] + sum[1] + sum[2]) / std::max(1, cnt[0] + cnt[1] + cnt[2]);
        for (int k = 0; k < 3; ++k) out[k] = (float) (cnt[k] ? sum[k] / cnt[k] : all);
        return true;
    }
    auto l = to_list(*p, p->get("value"));
    if (l.size() == 1) out[0] = out[1] = out[2] = (float) l[0];
    else if (l.size() == 3) for (int i = 0; i < 3; ++i) out[i] = (float) l[(size_t) i];
    else fail("Error while loading \"" + n.file + "\": \"" + name + "\" needs 1 or 3 components");
    return true;
}

}  // namespace
}  // namespace vrlhost

#include "vrl_xml_scene.inl"
