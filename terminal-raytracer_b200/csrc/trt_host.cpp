// trt_host.cpp -- native restatement of the reference's host logic around the hot path (no CUDA
// here): camera (src/camera.rs), scene file schema and primitive flattening (src/lib.rs:52-154,
// src/primitive.rs, src/vec3.rs), per-frame uniforms (src/lib.rs:417-439), key semantics
// (src/lib.rs:392-405).  Citations are relative to /root/reference/.  Compiled with
// -ffp-contract=off: every f32 expression rounds once per operator like the Rust it mirrors.
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/trt_b200.h"

// ---- camera (camera.rs:4-26, lib.rs:118, lib.rs:392-405) ----------------------------------------------
extern "C" void trt_camera_default(trt_camera *cam) {
    if (!cam) return;
    cam->position = trt_vec3{0.0f, 0.0f, 0.0f, 0.0f};
    cam->yaw = -3.14159274101257324f / 2.0f;      // -std::f32::consts::PI / 2.0
    cam->pitch = 0.0f;
}

extern "C" void trt_camera_vectors(const trt_camera *cam, trt_vec3 *forward, trt_vec3 *right, trt_vec3 *up) {
    if (!cam) return;
    // f32::cos / f32::sin are libm cosf / sinf on Linux
    trt_vec3 f{cosf(cam->yaw) * cosf(cam->pitch), sinf(cam->pitch), sinf(cam->yaw) * cosf(cam->pitch), 0.0f};
    trt_vec3 r{-sinf(cam->yaw), 0.0f, cosf(cam->yaw), 0.0f};
    trt_vec3 u{r.y * f.z - r.z * f.y, r.z * f.x - r.x * f.z, r.x * f.y - r.y * f.x, 0.0f};   // right.cross(forward), vec3.rs:23-29
    if (forward) *forward = f;
    if (right) *right = r;
    if (up) *up = u;
}

extern "C" int trt_camera_key(trt_camera *cam, int key) {
    if (!cam) return 0;
    trt_vec3 f, r;
    trt_camera_vectors(cam, &f, &r, nullptr);
    int moved = 1;
    trt_vec3 &p = cam->position;
    switch (key) {
        case 'w': p = trt_vec3{p.x + f.x * 0.1f, p.y + f.y * 0.1f, p.z + f.z * 0.1f, 0.0f}; break;
        case 's': p = trt_vec3{p.x - f.x * 0.1f, p.y - f.y * 0.1f, p.z - f.z * 0.1f, 0.0f}; break;
        case 'a': p = trt_vec3{p.x - r.x * 0.1f, p.y - r.y * 0.1f, p.z - r.z * 0.1f, 0.0f}; break;
        case 'd': p = trt_vec3{p.x + r.x * 0.1f, p.y + r.y * 0.1f, p.z + r.z * 0.1f, 0.0f}; break;
        case TRT_KEY_UP: cam->pitch += 0.05f; break;
        case TRT_KEY_DOWN: cam->pitch -= 0.05f; break;
        case TRT_KEY_LEFT: cam->yaw -= 0.05f; break;
        case TRT_KEY_RIGHT: cam->yaw += 0.05f; break;
        default: moved = 0; break;
    }
    // lib.rs:405 runs for every key event: f32::clamp(-1.5, 1.5)
    if (cam->pitch < -1.5f) cam->pitch = -1.5f;
    if (cam->pitch > 1.5f) cam->pitch = 1.5f;
    return moved;
}

// ---- a small JSON reader (objects, arrays, numbers, strings, true/false/null) ---------------------------
namespace {

struct JValue {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    double num = 0.0;
    bool is_int = false;          // written without fraction/exponent
    bool boolean = false;
    std::string str;
    std::vector<JValue> arr;
    std::vector<std::pair<std::string, JValue>> obj;
    const JValue *get(const char *key) const {
        const JValue *found = nullptr;                 // serde keeps the last duplicate
        for (auto &kv : obj) if (kv.first == key) found = &kv.second;
        return found;
    }
};

struct JParser {
    const char *p, *end;
    std::string err;
    int depth = 0;
    explicit JParser(const char *s, size_t n) : p(s), end(s + n) {}
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool fail(const std::string &m) { if (err.empty()) err = m; return false; }
    bool value(JValue &v) {
        if (++depth > 128) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end of input");
        bool ok;
        char c = *p;
        if (c == '{') ok = object(v);
        else if (c == '[') ok = array(v);
        else if (c == '"') { v.kind = JValue::Str; ok = string(v.str); }
        else if (c == '-' || (c >= '0' && c <= '9')) ok = number(v);
        else if (lit("true")) { v.kind = JValue::Bool; v.boolean = true; ok = true; }
        else if (lit("false")) { v.kind = JValue::Bool; ok = true; }
        else if (lit("null")) { v.kind = JValue::Null; ok = true; }
        else ok = fail(std::string("unexpected character '") + c + "'");
        depth--;
        return ok;
    }
    bool lit(const char *s) {
        size_t n = strlen(s);
        if ((size_t)(end - p) >= n && memcmp(p, s, n) == 0) { p += n; return true; }
        return false;
    }
    bool number(JValue &v) {
        const char *s = p;
        if (p < end && *p == '-') p++;
        if (p >= end || !isdigit((unsigned char)*p)) return fail("bad number");
        if (*p == '0') p++; else while (p < end && isdigit((unsigned char)*p)) p++;
        bool integral = true;
        if (p < end && *p == '.') { integral = false; p++; if (p >= end || !isdigit((unsigned char)*p)) return fail("bad number"); while (p < end && isdigit((unsigned char)*p)) p++; }
        if (p < end && (*p == 'e' || *p == 'E')) {
            integral = false; p++;
            if (p < end && (*p == '+' || *p == '-')) p++;
            if (p >= end || !isdigit((unsigned char)*p)) return fail("bad number");
            while (p < end && isdigit((unsigned char)*p)) p++;
        }
        std::string tmp(s, p);
        v.kind = JValue::Num;
        v.num = strtod(tmp.c_str(), nullptr);        // correctly rounded f64, as serde_json
        v.is_int = integral;
        return true;
    }
    bool string(std::string &out) {
        p++;   // opening quote
        out.clear();
        while (p < end && *p != '"') {
            if (*p == '\\') {
                p++;
                if (p >= end) return fail("bad escape");
                switch (*p) {
                    case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                    case 'u': if (end - p < 5) return fail("bad \\u escape"); out += '?'; p += 4; break;
                    default: out += *p; break;
                }
                p++;
            } else out += *p++;
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool array(JValue &v) {
        v.kind = JValue::Arr;
        p++;
        ws();
        if (p < end && *p == ']') { p++; return true; }
        for (;;) {
            v.arr.emplace_back();
            if (!value(v.arr.back())) return false;
            ws();
            if (p < end && *p == ',') { p++; continue; }
            if (p < end && *p == ']') { p++; return true; }
            return fail("expected ',' or ']'");
        }
    }
    bool object(JValue &v) {
        v.kind = JValue::Obj;
        p++;
        ws();
        if (p < end && *p == '}') { p++; return true; }
        for (;;) {
            ws();
            if (p >= end || *p != '"') return fail("expected a string key");
            std::string key;
            if (!string(key)) return false;
            ws();
            if (p >= end || *p != ':') return fail("expected ':'");
            p++;
            v.obj.emplace_back(key, JValue());
            if (!value(v.obj.back().second)) return false;
            ws();
            if (p < end && *p == ',') { p++; continue; }
            if (p < end && *p == '}') { p++; return true; }
            return fail("expected ',' or '}'");
        }
    }
};

}  // namespace

// ---- scene (lib.rs:52-98 schema, lib.rs:120-154 flattening) --------------------------------------------
struct trt_scene {
    trt_scene_info info{};
    std::vector<trt_primitive> prims;
};

namespace {

// the reference's embedded default scene (lib.rs:107 include_str!("scenes/Cornell_Box.json")), as data
const char *kCornellBox = R"JSON({
 "width": 400, "height": 200, "samples_per_pixel": 128, "max_depth": 3, "frames_to_accumulate": 100000000,
 "camera": {"fov_degrees": 45.0, "char_aspect_ratio": 0.55},
 "spheres": [
  {"center": [-0.5, -0.4, -2.5], "radius": 0.6, "color": [0.9, 0.9, 0.9], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.5},
  {"center": [0.5, -0.7, -3.2], "radius": 0.3, "color": [0.6, 0.8, 0.9], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"center": [0.0, 0.1, -3.2], "radius": 0.4, "color": [0.1, 0.2, 0.4], "emission": [0.0, 0.0, 0.0], "reflectivity": 1.0}],
 "planes": [
  {"point": [0.0, -1.0, 0.0], "normal": [0.0, 1.0, 0.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"point": [0.0, 1.0, 0.0], "normal": [0.0, -1.0, 0.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"point": [1.0, 0.0, 0.0], "normal": [-1.0, 0.0, 0.0], "color": [0.25, 0.75, 0.25], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"point": [-1.0, 0.0, 0.0], "normal": [1.0, 0.0, 0.0], "color": [0.75, 0.25, 0.25], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"point": [0.0, 0.0, 1.0], "normal": [0.0, 0.0, -1.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
  {"point": [0.0, 0.0, -4.0], "normal": [0.0, 0.0, 1.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0}],
 "triangles": [
  {"v0": [-0.4, 0.99, -2.4], "v1": [0.4, 0.99, -2.4], "v2": [-0.4, 0.99, -3.2], "color": [1.0, 1.0, 1.0], "emission": [8.0, 8.0, 8.0], "reflectivity": 0.0},
  {"v0": [0.4, 0.99, -2.4], "v1": [0.4, 0.99, -3.2], "v2": [-0.4, 0.99, -3.2], "color": [1.0, 1.0, 1.0], "emission": [8.0, 8.0, 8.0], "reflectivity": 0.0}]
})JSON";

struct Err {
    std::string msg;
    bool set(const std::string &m) { if (msg.empty()) msg = m; return false; }
};

bool get_u32(const JValue &o, const char *key, uint32_t &out, Err &e) {
    const JValue *v = o.get(key);
    if (!v) return e.set(std::string("missing field `") + key + "`");
    if (v->kind != JValue::Num || !v->is_int || v->num < 0.0 || v->num > 4294967295.0)
        return e.set(std::string("field `") + key + "`: expected u32");
    out = (uint32_t)v->num;
    return true;
}
bool get_f64(const JValue &o, const char *key, double &out, Err &e) {
    const JValue *v = o.get(key);
    if (!v) return e.set(std::string("missing field `") + key + "`");
    if (v->kind != JValue::Num) return e.set(std::string("field `") + key + "`: expected a number");
    out = v->num;
    return true;
}
// [f64; 3] -> Vec3::new(x, y, z): `as f32` casts (vec3.rs:15-17)
bool get_vec3(const JValue &o, const char *key, trt_vec3 &out, Err &e) {
    const JValue *v = o.get(key);
    if (!v) return e.set(std::string("missing field `") + key + "`");
    if (v->kind != JValue::Arr || v->arr.size() != 3) return e.set(std::string("field `") + key + "`: expected an array of 3 numbers");
    for (auto &c : v->arr) if (c.kind != JValue::Num) return e.set(std::string("field `") + key + "`: expected an array of 3 numbers");
    out = trt_vec3{(float)v->arr[0].num, (float)v->arr[1].num, (float)v->arr[2].num, 0.0f};
    return true;
}
bool material(const JValue &o, trt_primitive &p, Err &e) {
    double refl;
    if (!get_vec3(o, "color", p.color, e) || !get_vec3(o, "emission", p.emission, e) || !get_f64(o, "reflectivity", refl, e)) return false;
    p.reflectivity = (float)refl;
    return true;
}

bool build_scene(const JValue &root, trt_scene &s, Err &e) {
    if (root.kind != JValue::Obj) return e.set("expected a JSON object");
    trt_scene_info &in = s.info;
    if (!get_u32(root, "width", in.width, e) || !get_u32(root, "height", in.height, e) ||
        !get_u32(root, "samples_per_pixel", in.samples_per_pixel, e) || !get_u32(root, "max_depth", in.max_depth, e) ||
        !get_u32(root, "frames_to_accumulate", in.frames_to_accumulate, e))
        return false;
    const JValue *cam = root.get("camera");
    if (!cam || cam->kind != JValue::Obj) return e.set("missing field `camera`");
    double fov, car;
    if (!get_f64(*cam, "fov_degrees", fov, e) || !get_f64(*cam, "char_aspect_ratio", car, e)) return false;
    in.fov_degrees = (float)fov;
    in.char_aspect_ratio = (float)car;
    const JValue *sp = root.get("spheres"), *pl = root.get("planes"), *tr = root.get("triangles");
    if (!sp || sp->kind != JValue::Arr) return e.set("missing field `spheres`");        // mandatory (lib.rs:60)
    if (!pl || pl->kind != JValue::Arr) return e.set("missing field `planes`");         // mandatory (lib.rs:61)
    if (tr && tr->kind != JValue::Arr) return e.set("field `triangles`: expected an array");   // #[serde(default)]
    for (auto &o : sp->arr) {                                                             // primitive.rs:37-54
        trt_primitive p;
        memset(&p, 0, sizeof p);
        double radius;
        if (o.kind != JValue::Obj) return e.set("spheres: expected objects");
        if (!get_vec3(o, "center", p.sphere_center, e) || !get_f64(o, "radius", radius, e) || !material(o, p, e)) return false;
        p.primitive_type = 0;
        p.sphere_radius = (float)radius;
        s.prims.push_back(p);
    }
    for (auto &o : pl->arr) {                                                             // primitive.rs:56-73
        trt_primitive p;
        memset(&p, 0, sizeof p);
        if (o.kind != JValue::Obj) return e.set("planes: expected objects");
        if (!get_vec3(o, "point", p.plane_point, e) || !get_vec3(o, "normal", p.plane_normal, e) || !material(o, p, e)) return false;
        p.primitive_type = 1;
        s.prims.push_back(p);
    }
    if (tr) for (auto &o : tr->arr) {                                                     // primitive.rs:75-92
        trt_primitive p;
        memset(&p, 0, sizeof p);
        if (o.kind != JValue::Obj) return e.set("triangles: expected objects");
        if (!get_vec3(o, "v0", p.triangle_v0, e) || !get_vec3(o, "v1", p.triangle_v1, e) || !get_vec3(o, "v2", p.triangle_v2, e) ||
            !material(o, p, e))
            return false;
        p.primitive_type = 2;
        s.prims.push_back(p);
    }
    in.n_spheres = (uint32_t)sp->arr.size();
    in.n_planes = (uint32_t)pl->arr.size();
    in.n_triangles = tr ? (uint32_t)tr->arr.size() : 0u;
    return true;
}

void put_err(char *err, size_t cap, const std::string &m) {
    if (err && cap) { snprintf(err, cap, "%s", m.c_str()); }
}

}  // namespace

static_assert(sizeof(trt_vec3) == 16, "Vec3 is 16 bytes (vec3.rs)");
static_assert(sizeof(trt_primitive) == 180, "Primitive is 180 bytes (primitive.rs)");
static_assert(sizeof(trt_uniforms) == 160, "Uniforms is 160 bytes (lib.rs:23-50)");
static_assert(sizeof(trt_variance) == 8, "VarianceBuffer is 8 bytes (shader.wgsl:71-74)");

extern "C" int trt_scene_parse(const char *json_text, size_t json_len, trt_scene **out, char *err, size_t err_cap) {
    if (!json_text || !out) return TRT_ERR_INVALID;
    *out = nullptr;
    JParser jp(json_text, json_len);
    JValue root;
    if (!jp.value(root)) { put_err(err, err_cap, "Failed to parse scene: " + jp.err); return TRT_ERR_PARSE; }
    jp.ws();
    if (jp.p != jp.end) { put_err(err, err_cap, "Failed to parse scene: trailing characters"); return TRT_ERR_PARSE; }
    std::unique_ptr<trt_scene> s(new (std::nothrow) trt_scene());
    if (!s) return TRT_ERR_NOMEM;
    Err e;
    if (!build_scene(root, *s, e)) { put_err(err, err_cap, "Failed to parse scene: " + e.msg); return TRT_ERR_PARSE; }
    *out = s.release();
    return TRT_OK;
}

extern "C" int trt_scene_load(const char *path, trt_scene **out, char *err, size_t err_cap) {
    if (!out) return TRT_ERR_INVALID;
    if (!path) return trt_scene_parse(kCornellBox, strlen(kCornellBox), out, err, err_cap);     // lib.rs:107
    FILE *f = fopen(path, "rb");
    if (!f) { put_err(err, err_cap, std::string("Failed to read scene file at '") + path + "'"); return TRT_ERR_IO; }   // lib.rs:105-106
    std::string text;
    char buf[65536];
    size_t n;
    while ((n = fread(buf, 1, sizeof buf, f)) > 0) text.append(buf, n);
    fclose(f);
    return trt_scene_parse(text.data(), text.size(), out, err, err_cap);
}

extern "C" void trt_scene_free(trt_scene *s) { delete s; }

extern "C" int trt_scene_get_info(const trt_scene *s, trt_scene_info *info) {
    if (!s || !info) return TRT_ERR_INVALID;
    *info = s->info;
    return TRT_OK;
}

extern "C" void trt_scene_clamp(trt_scene *s, uint32_t cols, uint32_t rows) {     // lib.rs:114-115
    if (!s) return;
    if (cols < s->info.width) s->info.width = cols;
    const uint32_t r2 = rows >= 2u ? rows - 2u : 0u;
    if (r2 < s->info.height) s->info.height = r2;
}

extern "C" void trt_scene_set_size(trt_scene *s, uint32_t w, uint32_t h) {
    if (!s) return;
    s->info.width = w;
    s->info.height = h;
}

extern "C" const trt_primitive *trt_scene_primitives(const trt_scene *s, uint32_t *count) {
    if (!s) { if (count) *count = 0; return nullptr; }
    if (count) *count = (uint32_t)s->prims.size();
    return s->prims.data();
}

extern "C" void trt_fill_uniforms(const trt_scene *s, const trt_camera *cam, uint32_t seed, uint32_t frame_number,
                                  trt_uniforms *out) {                           // lib.rs:417-439
    if (!s || !cam || !out) return;
    memset(out, 0, sizeof *out);                                                 // grid fields: dead, zero
    trt_vec3 f, r, u;
    trt_camera_vectors(cam, &f, &r, &u);
    out->width = s->info.width;
    out->height = s->info.height;
    out->samples_per_pixel = s->info.samples_per_pixel;
    out->max_depth = s->info.max_depth;
    out->seed = seed;
    out->frame_number = frame_number;
    out->primitive_count = (uint32_t)s->prims.size();
    out->aspect_ratio = (float)s->info.width / (float)s->info.height;
    out->char_aspect_ratio = s->info.char_aspect_ratio;
    out->fov_rad = s->info.fov_degrees * (3.14159274101257324f / 180.0f);        // f32::to_radians
    out->camera_pos = cam->position;
    out->camera_forward = f;
    out->camera_right = r;
    out->camera_up = u;
}
