// sba_project.cpp -- reader / writer of the reference's project files (".steviapproj", JSON) and their conversion to the flat
// arrays of include/sba_b200.h (SURVEY 8f row 4, I/O half).  Host code only, no CUDA, no Qt: a small JSON document model
// stands in for QJsonDocument.
//
// What is mirrored (reference files under /root/reference/libs):
//   file layout       datablocks/project.cpp:93-221   one array per data-block class name + "metadata" {local_transform {R, t}}
//   block envelope    datablocks/project.cpp:1219-1288  id, name, "fixed" / "unfixed", "enabled" / "disabled"
//   floatParameter    datablocks/floatparameter.cpp:256-296  {isSet, val, isUncertain, stddev}; group: {isSet, vals[n]}
//   Landmark          datablocks/point3d.cpp:155-184  x y z op (+ geo_crs, landmark.cpp:296-314)
//   Image             datablocks/rigidbody.cpp:344-392 x y z rx ry rz op or; image.cpp:368-443 assignedCamera, assignedTrajectory,
//                     assignedMounting, Landmarks[] {attachedLandmarkId, x, y} (image.cpp:715-737)
//   Camera            datablocks/camera.cpp:548-700   f cx cy k1..k3 p1 p2 B1 B2 useRd useTd useSd and the optimised "o*" twins
// and how the SBA modules turn the blocks into a problem:
//   which items take part   sparsesolver/sbagraphreductor.cpp:172-194 (drop every node with fewer observations than degrees
//                           of freedom, to the fixed point); node / observation counts sbamodules/landmarkssbamodule.cpp:21-74,
//                           sbamodules/imagealignementsbamodule.cpp:30-112
//   landmark values/priors  sbamodules/landmarkssbamodule.cpp:111-205 (optimised position, else the prior, else 0; prior
//                           A = diag(1/|stddev| or 1e6); fixed -> constant)
//   pose values / priors    sparsesolver/modularsbasolver.cpp:233-305, sbamodules/imagealignementsbamodule.cpp:206-286
//   observations            sbamodules/imagealignementsbamodule.cpp:388-440 (info = diag(1/stddev or 1))
//   intrinsic blocks/priors sbamodules/pinholecameraprojectormodule.cpp:302-533
//   write-back              landmarkssbamodule.cpp:207-262, imagealignementsbamodule.cpp:446-490,
//                           pinholecameraprojectormodule.cpp:535-574 (values pass through float, as there)
// Not supported (sba_project_open / _convert fail with the reason): landmarks with a geographic CRS (needs PROJ), images
// attached to a trajectory + mounting (trajectory factors are outside the engine).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <string>
#include <utility>
#include <vector>

#include "sba_b200.h"

static thread_local std::string g_proj_error;
static int proj_fail(const std::string& m) { g_proj_error = m; return SBA_ERR_INVALID; }

namespace {

// ---------------------------------------------------------------------------
// JSON document model
// ---------------------------------------------------------------------------
struct JVal {
    enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;   // file order; written sorted by key (QJsonObject keeps its keys sorted)

    const JVal* get(const char* key) const {
        if (type != Obj) return nullptr;
        for (const auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    JVal* get(const char* key) { return const_cast<JVal*>(static_cast<const JVal*>(this)->get(key)); }
    JVal& set(const char* key, JVal v) {
        if (JVal* e = get(key)) { *e = std::move(v); return *e; }
        obj.emplace_back(key, std::move(v));
        return obj.back().second;
    }
    // QJsonValue::toDouble / toInt / toBool / toString with their defaults
    double to_double(double def = 0.0) const { return type == Num ? num : def; }
    long long to_int(long long def = 0) const { return (type == Num && num == std::floor(num)) ? (long long)num : def; }
    bool to_bool() const { return type == Bool && b; }
    static JVal number(double v) { JVal j; j.type = Num; j.num = v; return j; }
    static JVal object() { JVal j; j.type = Obj; return j; }
    static JVal array() { JVal j; j.type = Arr; return j; }
};

struct Parser {
    const char* p;
    const char* end;
    std::string err;
    int depth = 0;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    static void utf8(std::string& s, unsigned cp) {
        if (cp < 0x80) s += (char)cp;
        else if (cp < 0x800) { s += (char)(0xC0 | cp >> 6); s += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { s += (char)(0xE0 | cp >> 12); s += (char)(0x80 | (cp >> 6 & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
        else { s += (char)(0xF0 | cp >> 18); s += (char)(0x80 | (cp >> 12 & 0x3F)); s += (char)(0x80 | (cp >> 6 & 0x3F)); s += (char)(0x80 | (cp & 0x3F)); }
    }
    bool hex4(unsigned& v) {
        if (end - p < 4) return fail("truncated \\u escape");
        v = 0;
        for (int i = 0; i < 4; i++) {
            const char c = *p++;
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else return fail("bad \\u escape");
        }
        return true;
    }
    bool string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected a string");
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("truncated escape");
                const char c = *p++;
                switch (c) {
                    case '"': out += '"'; break;
                    case '\\': out += '\\'; break;
                    case '/': out += '/'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'n': out += '\n'; break;
                    case 'r': out += '\r'; break;
                    case 't': out += '\t'; break;
                    case 'u': {
                        unsigned cp;
                        if (!hex4(cp)) return false;
                        if (cp >= 0xD800 && cp < 0xDC00 && end - p >= 6 && p[0] == '\\' && p[1] == 'u') {   // surrogate pair
                            p += 2;
                            unsigned lo;
                            if (!hex4(lo)) return false;
                            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                        }
                        utf8(out, cp);
                        break;
                    }
                    default: return fail("unknown escape");
                }
            } else {
                out += *p++;
            }
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool value(JVal& v) {
        if (++depth > 256) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end of file");
        bool ok = true;
        if (*p == '{') {
            p++;
            v.type = JVal::Obj;
            ws();
            if (p < end && *p == '}') { p++; }
            else
                for (;;) {
                    ws();
                    std::string key;
                    if (!string(key)) { ok = false; break; }
                    ws();
                    if (p >= end || *p != ':') { ok = fail("expected ':'"); break; }
                    p++;
                    v.obj.emplace_back(std::move(key), JVal());
                    if (!value(v.obj.back().second)) { ok = false; break; }
                    ws();
                    if (p < end && *p == ',') { p++; continue; }
                    if (p < end && *p == '}') { p++; break; }
                    ok = fail("expected ',' or '}'");
                    break;
                }
        } else if (*p == '[') {
            p++;
            v.type = JVal::Arr;
            ws();
            if (p < end && *p == ']') { p++; }
            else
                for (;;) {
                    v.arr.emplace_back();
                    if (!value(v.arr.back())) { ok = false; break; }
                    ws();
                    if (p < end && *p == ',') { p++; continue; }
                    if (p < end && *p == ']') { p++; break; }
                    ok = fail("expected ',' or ']'");
                    break;
                }
        } else if (*p == '"') {
            v.type = JVal::Str;
            ok = string(v.str);
        } else if (end - p >= 4 && !std::strncmp(p, "true", 4)) { v.type = JVal::Bool; v.b = true; p += 4; }
        else if (end - p >= 5 && !std::strncmp(p, "false", 5)) { v.type = JVal::Bool; v.b = false; p += 5; }
        else if (end - p >= 4 && !std::strncmp(p, "null", 4)) { v.type = JVal::Null; p += 4; }
        else {
            char* e = nullptr;
            const double d = std::strtod(p, &e);   // the buffer is NUL-terminated
            if (e == p) ok = fail("unexpected character");
            else { v.type = JVal::Num; v.num = d; p = e; }
        }
        depth--;
        return ok;
    }
};

void write_string(std::string& out, const std::string& s) {
    out += '"';
    for (const unsigned char c : s) {
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\b': out += "\\b"; break;
            case '\f': out += "\\f"; break;
            case '\n': out += "\\n"; break;
            case '\r': out += "\\r"; break;
            case '\t': out += "\\t"; break;
            default:
                if (c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); out += b; }
                else out += (char)c;
        }
    }
    out += '"';
}
void write_number(std::string& out, double v) {
    char b[40];
    if (!std::isfinite(v)) { out += "null"; return; }   // as QJsonDocument writes a non-finite double
    if (v == std::floor(v) && std::fabs(v) < 9007199254740992.0) { std::snprintf(b, sizeof b, "%.0f", v); out += b; return; }
    for (int prec = 15; prec <= 17; prec++) {   // shortest form that reads back bit-exact
        std::snprintf(b, sizeof b, "%.*g", prec, v);
        if (std::strtod(b, nullptr) == v) break;
    }
    out += b;
}
void write_value(std::string& out, const JVal& v, int indent) {
    const std::string pad((size_t)indent * 4, ' '), pad1((size_t)(indent + 1) * 4, ' ');
    switch (v.type) {
        case JVal::Null: out += "null"; break;
        case JVal::Bool: out += v.b ? "true" : "false"; break;
        case JVal::Num: write_number(out, v.num); break;
        case JVal::Str: write_string(out, v.str); break;
        case JVal::Arr:
            if (v.arr.empty()) { out += "[\n" + pad + "]"; break; }
            out += "[\n";
            for (size_t i = 0; i < v.arr.size(); i++) {
                out += pad1;
                write_value(out, v.arr[i], indent + 1);
                out += i + 1 < v.arr.size() ? ",\n" : "\n";
            }
            out += pad + "]";
            break;
        case JVal::Obj: {
            if (v.obj.empty()) { out += "{\n" + pad + "}"; break; }
            std::vector<const std::pair<std::string, JVal>*> order;
            for (const auto& kv : v.obj) order.push_back(&kv);
            std::stable_sort(order.begin(), order.end(), [](const auto* a, const auto* b) { return a->first < b->first; });
            out += "{\n";
            for (size_t i = 0; i < order.size(); i++) {
                out += pad1;
                write_string(out, order[i]->first);
                out += ": ";
                write_value(out, order[i]->second, indent + 1);
                out += i + 1 < order.size() ? ",\n" : "\n";
            }
            out += pad + "}";
            break;
        }
    }
}

// ---------------------------------------------------------------------------
// data blocks
// ---------------------------------------------------------------------------
struct FloatParam {   // floatParameter::fromJson (floatparameter.cpp:267-296)
    bool is_set = false, is_uncertain = false;
    double value = 0.0, stddev = 1.0;
};
FloatParam float_param(const JVal* o) {
    FloatParam f;
    if (!o || o->type != JVal::Obj) return f;
    if (const JVal* v = o->get("val")) f.value = v->to_double();
    if (const JVal* v = o->get("stddev")) f.stddev = v->to_double(1.0);
    if (const JVal* v = o->get("isSet")) f.is_set = v->to_bool() || v->to_int() > 0;
    if (const JVal* v = o->get("isUncertain")) f.is_uncertain = v->to_int() != 0;
    return f;
}
struct FloatGroup3 {   // floatParameterGroup<3>::fromJson (floatparameter.h:251-272): values only when exactly 3 are stored
    bool is_set = false;
    double v[3] = {0.0, 0.0, 0.0};
};
FloatGroup3 float_group3(const JVal* o) {
    FloatGroup3 g;
    if (!o || o->type != JVal::Obj) return g;
    const JVal* vals = o->get("vals");
    if (vals && vals->type == JVal::Arr && vals->arr.size() == 3) {
        for (int i = 0; i < 3; i++) g.v[i] = vals->arr[i].to_double();
        if (const JVal* s = o->get("isSet")) g.is_set = s->to_bool() || s->to_int() > 0;
    }
    return g;
}
JVal float_param_json(double value) {   // floatParameter::toJson of a set, certain parameter (setOptimized*(float))
    JVal o = JVal::object();
    o.set("isSet", JVal::number(1)); o.set("val", JVal::number(value));
    o.set("isUncertain", JVal::number(0)); o.set("stddev", JVal::number(1.0));
    return o;
}
JVal float_param_unset_json() {   // a default-constructed floatParameter (Camera::clearOptimized)
    JVal o = JVal::object();
    o.set("isSet", JVal::number(0)); o.set("val", JVal::number(0.0));
    o.set("isUncertain", JVal::number(0)); o.set("stddev", JVal::number(1.0));
    return o;
}
JVal float_group3_json(const double* v) {   // floatParameterGroup<3>::toJson (floatparameter.h:227-249; its CovMatTrig array is written empty there too)
    JVal o = JVal::object();
    o.set("isSet", JVal::number(1));
    JVal vals = JVal::array();
    for (int i = 0; i < 3; i++) vals.arr.push_back(JVal::number(v[i]));
    o.set("vals", std::move(vals));
    o.set("isUncertain", JVal::number(0));
    o.set("CovMatTrig", JVal::array());
    return o;
}

struct Block {   // DataBlock::setFromJson (project.cpp:1238-1288)
    long long id = -1;
    bool fixed = false, enabled = true;
    size_t json_index = 0;
};
std::string lower_trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && std::isspace((unsigned char)s[a])) a++;
    while (b > a && std::isspace((unsigned char)s[b - 1])) b--;
    std::string r = s.substr(a, b - a);
    for (char& c : r) c = (char)std::tolower((unsigned char)c);
    return r;
}
Block block_header(const JVal& o, size_t index) {
    Block b;
    b.json_index = index;
    if (const JVal* v = o.get("id")) b.id = v->to_int(-1);
    if (const JVal* v = o.get("fixed")) b.fixed = v->type == JVal::Str && lower_trim(v->str) == "fixed";
    if (const JVal* v = o.get("enabled")) b.enabled = v->type == JVal::Str && lower_trim(v->str) == "enabled";
    return b;
}

struct LandmarkBlock : Block {
    FloatParam x, y, z;
    FloatGroup3 op;
    bool geo = false;
};
struct ImageLandmarkRec {
    long long landmark_id = -1;
    FloatParam x, y;
};
struct ImageBlock : Block {
    FloatParam x, y, z, rx, ry, rz;
    FloatGroup3 op, orot;
    long long camera = -1, trajectory = -1, mounting = -1;
    std::vector<ImageLandmarkRec> lms;
};
struct CameraBlock : Block {
    FloatParam f, cx, cy, k1, k2, k3, p1, p2, B1, B2, of, ocx, ocy, ok1, ok2, ok3, op1, op2, oB1, oB2;
    bool use_rd = true, use_td = true, use_sd = true;   // Camera's constructor defaults (camera.cpp)
};

const char* const kLandmarkClass = "StereoVisionApp::Landmark";
const char* const kImageClass = "StereoVisionApp::Image";
const char* const kCameraClass = "StereoVisionApp::Camera";

}  // namespace

struct sba_project {
    JVal doc;
    std::vector<LandmarkBlock> landmarks;
    std::vector<ImageBlock> images;
    std::vector<CameraBlock> cameras;
    bool has_local = false;
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 0};
    // the problem (built by open): indices into the vectors above
    std::vector<int32_t> kept_images, kept_landmarks, kept_cameras;   // engine index -> block index
    std::vector<int32_t> obs_cam, obs_pt;
    std::vector<double> obs_uv, obs_info;
    int32_t n_priors = 0, n_intr_priors = 0;
    int32_t skipped_images_without_camera = 0;
};

namespace {

void to_local(const sba_project& p, const double* x, double* out) {   // getTransform2LocalFrame() * x (modularsbasolver.cpp:530-541)
    for (int i = 0; i < 3; i++) out[i] = p.R[3 * i] * x[0] + p.R[3 * i + 1] * x[1] + p.R[3 * i + 2] * x[2] + p.t[i];
}

int parse_blocks(sba_project& p) {
    const JVal& d = p.doc;
    if (const JVal* arr = d.get(kLandmarkClass))
        for (size_t i = 0; arr->type == JVal::Arr && i < arr->arr.size(); i++) {
            const JVal& o = arr->arr[i];
            LandmarkBlock b;
            static_cast<Block&>(b) = block_header(o, i);
            if (b.id < 0) continue;   // Project::load keeps only blocks with an id (project.cpp:131-137)
            b.x = float_param(o.get("x")); b.y = float_param(o.get("y"));
            if (o.get("y")) b.z = float_param(o.get("z"));   // the reference tests "y" before reading "z" (point3d.cpp:176-178)
            b.op = float_group3(o.get("op"));
            if (const JVal* g = o.get("geo_crs")) b.geo = g->type == JVal::Str && !g->str.empty();
            p.landmarks.push_back(b);
        }
    if (const JVal* arr = d.get(kCameraClass))
        for (size_t i = 0; arr->type == JVal::Arr && i < arr->arr.size(); i++) {
            const JVal& o = arr->arr[i];
            CameraBlock b;
            static_cast<Block&>(b) = block_header(o, i);
            if (b.id < 0) continue;
            b.f = float_param(o.get("f")); b.cx = float_param(o.get("cx")); b.cy = float_param(o.get("cy"));
            b.k1 = float_param(o.get("k1")); b.k2 = float_param(o.get("k2")); b.k3 = float_param(o.get("k3"));
            b.p1 = float_param(o.get("p1")); b.p2 = float_param(o.get("p2"));
            b.B1 = float_param(o.get("B1")); b.B2 = float_param(o.get("B2"));
            b.of = float_param(o.get("of")); b.ocx = float_param(o.get("ocx")); b.ocy = float_param(o.get("ocy"));
            b.ok1 = float_param(o.get("ok1")); b.ok2 = float_param(o.get("ok2")); b.ok3 = float_param(o.get("ok3"));
            b.op1 = float_param(o.get("op1")); b.op2 = float_param(o.get("op2"));
            b.oB1 = float_param(o.get("oB1")); b.oB2 = float_param(o.get("oB2"));
            if (const JVal* v = o.get("useRd")) b.use_rd = v->to_bool();
            if (const JVal* v = o.get("useTd")) b.use_td = v->to_bool();
            if (const JVal* v = o.get("useSd")) b.use_sd = v->to_bool();
            p.cameras.push_back(b);
        }
    if (const JVal* arr = d.get(kImageClass))
        for (size_t i = 0; arr->type == JVal::Arr && i < arr->arr.size(); i++) {
            const JVal& o = arr->arr[i];
            ImageBlock b;
            static_cast<Block&>(b) = block_header(o, i);
            if (b.id < 0) continue;
            b.x = float_param(o.get("x")); b.y = float_param(o.get("y")); b.z = float_param(o.get("z"));
            b.rx = float_param(o.get("rx")); b.ry = float_param(o.get("ry")); b.rz = float_param(o.get("rz"));
            b.op = float_group3(o.get("op")); b.orot = float_group3(o.get("or"));
            if (const JVal* v = o.get("assignedCamera")) b.camera = v->to_int(0);
            if (const JVal* v = o.get("assignedTrajectory")) b.trajectory = v->to_int(0);
            if (const JVal* v = o.get("assignedMounting")) b.mounting = v->to_int(0);
            if (const JVal* l = o.get("Landmarks"))
                for (size_t k = 0; l->type == JVal::Arr && k < l->arr.size(); k++) {
                    const JVal& lo = l->arr[k];
                    const JVal* idv = lo.get("id");
                    if (!idv || idv->to_int(-1) < 0) continue;   // sub-items without an id are dropped (image.cpp:436-438)
                    ImageLandmarkRec r;
                    r.x = float_param(lo.get("x")); r.y = float_param(lo.get("y"));
                    if (const JVal* a = lo.get("attachedLandmarkId")) r.landmark_id = a->to_int(-1);
                    b.lms.push_back(r);
                }
            p.images.push_back(b);
        }
    if (const JVal* md = d.get("metadata"))
        if (const JVal* lt = md->get("local_transform")) {
            const JVal* R = lt->get("R");
            const JVal* t = lt->get("t");
            if (R && t && R->type == JVal::Arr && t->type == JVal::Arr && R->arr.size() >= 3 && t->arr.size() >= 3) {
                for (int i = 0; i < 3; i++) {
                    const JVal& row = R->arr[i];
                    for (int j = 0; j < 3; j++) p.R[3 * i + j] = (row.type == JVal::Arr && row.arr.size() > (size_t)j) ? row.arr[j].to_double() : 0.0;
                    p.t[i] = t->arr[i].to_double();
                }
                p.has_local = true;
            }
        }
    return SBA_OK;
}

// GenericSBAGraphReductor (sbagraphreductor.cpp:172-308) on images (6 dof) and landmarks (3 dof)
int build_problem(sba_project& p) {
    std::map<long long, int32_t> lm_of, cam_of;
    for (size_t i = 0; i < p.landmarks.size(); i++) lm_of.emplace(p.landmarks[i].id, (int32_t)i);   // first block of an id wins, as QMap::insert in load order would be overwritten -- ids are unique in files the reference writes
    for (size_t i = 0; i < p.cameras.size(); i++) cam_of.emplace(p.cameras[i].id, (int32_t)i);
    const size_t NI = p.images.size(), NL = p.landmarks.size();
    for (size_t l = 0; l < NL; l++)
        if (p.landmarks[l].enabled && p.landmarks[l].geo)
            return proj_fail("landmark " + std::to_string(p.landmarks[l].id) + " carries a geographic CRS (geo_crs): conversion needs PROJ, not supported");
    std::vector<int> im_dof_obs(NI, 0), lm_dof_obs(NL, 0);
    std::vector<uint8_t> im_in(NI, 1), lm_in(NL, 0);
    for (size_t l = 0; l < NL; l++) lm_in[l] = p.landmarks[l].enabled ? 1 : 0;   // landmarkssbamodule.cpp:27-42
    for (size_t l = 0; l < NL; l++)
        if (lm_in[l] && p.landmarks[l].x.is_set && p.landmarks[l].y.is_set && p.landmarks[l].z.is_set) lm_dof_obs[l] += 3;
    // pair observation counts (image, landmark) -> 2 per attached landmark
    std::vector<std::vector<std::pair<int32_t, int>>> im_edges(NI);
    for (size_t i = 0; i < NI; i++) {
        const ImageBlock& im = p.images[i];
        if (!im.enabled) continue;
        if (im.trajectory >= 0 && im.mounting >= 0)   // imagealignementsbamodule.cpp:63-68,142-173
            return proj_fail("image " + std::to_string(im.id) + " is attached to a trajectory and a mounting: trajectory factors are outside the GPU engine");
        int self = 0;
        if (im.x.is_set && im.y.is_set && im.z.is_set) self += 3;
        if (im.rx.is_set && im.ry.is_set && im.rz.is_set) self += 3;
        im_dof_obs[i] += self;
        std::map<int32_t, int> per_lm;
        for (const ImageLandmarkRec& r : im.lms) {
            auto it = lm_of.find(r.landmark_id);
            if (it == lm_of.end() || !lm_in[it->second]) continue;
            per_lm[it->second] += 2;
        }
        for (const auto& e : per_lm) {
            im_edges[i].push_back(e);
            im_dof_obs[i] += e.second;
            lm_dof_obs[e.first] += e.second;
        }
    }
    std::vector<std::vector<std::pair<int32_t, int>>> lm_edges(NL);
    for (size_t i = 0; i < NI; i++)
        for (const auto& e : im_edges[i]) lm_edges[e.first].push_back({(int32_t)i, e.second});
    for (bool again = true; again;) {
        again = false;
        for (size_t i = 0; i < NI; i++)
            if (im_in[i] && im_dof_obs[i] < 6) {
                im_in[i] = 0; again = true;
                for (const auto& e : im_edges[i]) if (lm_in[e.first]) lm_dof_obs[e.first] -= e.second;
            }
        for (size_t l = 0; l < NL; l++)
            if (lm_in[l] && lm_dof_obs[l] < 3) {
                lm_in[l] = 0; again = true;
                for (const auto& e : lm_edges[l]) if (im_in[e.first]) im_dof_obs[e.first] -= e.second;
            }
    }
    // kept items, in file order; the camera models in order of first use (imagealignementsbamodule.cpp:177-196)
    std::vector<int32_t> lm_engine(NL, -1), cam_engine(p.cameras.size(), -1);
    for (size_t l = 0; l < NL; l++)
        if (lm_in[l]) { lm_engine[l] = (int32_t)p.kept_landmarks.size(); p.kept_landmarks.push_back((int32_t)l); }
    for (size_t i = 0; i < NI; i++) {
        if (!im_in[i]) continue;
        const ImageBlock& im = p.images[i];
        auto ci = cam_of.find(im.camera);
        if (ci == cam_of.end()) { p.skipped_images_without_camera++; continue; }   // no projector: the frame adds no projection cost
        if (cam_engine[ci->second] < 0) { cam_engine[ci->second] = (int32_t)p.kept_cameras.size(); p.kept_cameras.push_back(ci->second); }
        const int32_t e = (int32_t)p.kept_images.size();
        p.kept_images.push_back((int32_t)i);
        for (const ImageLandmarkRec& r : im.lms) {
            auto it = lm_of.find(r.landmark_id);
            if (it == lm_of.end() || !p.landmarks[it->second].enabled || lm_engine[it->second] < 0) continue;
            p.obs_cam.push_back(e); p.obs_pt.push_back(lm_engine[it->second]);
            p.obs_uv.push_back(r.x.value); p.obs_uv.push_back(r.y.value);
            p.obs_info.push_back(r.x.is_uncertain ? 1.0 / r.x.stddev : 1.0); p.obs_info.push_back(0.0);
            p.obs_info.push_back(0.0); p.obs_info.push_back(r.y.is_uncertain ? 1.0 / r.y.stddev : 1.0);
        }
    }
    return SBA_OK;
}

// the priors of the kept items; n only counts when the outputs are NULL
int32_t emit_priors(const sba_project& p, int32_t* kind, int32_t* index, double* A9, double* b3) {
    int32_t n = 0;
    auto put = [&](int32_t k, int32_t idx, const double* diag, const double* mean) {
        if (kind) {
            kind[n] = k; index[n] = idx;
            for (int i = 0; i < 9; i++) A9[9 * (size_t)n + i] = 0.0;
            for (int i = 0; i < 3; i++) { A9[9 * (size_t)n + 4 * i] = diag[i]; b3[3 * (size_t)n + i] = mean[i]; }
        }
        n++;
    };
    for (size_t e = 0; e < p.kept_landmarks.size(); e++) {   // landmarkssbamodule.cpp:167-201
        const LandmarkBlock& lm = p.landmarks[p.kept_landmarks[e]];
        if (lm.fixed || !(lm.x.is_set && lm.y.is_set && lm.z.is_set)) continue;
        const double raw[3] = {lm.x.value, lm.y.value, lm.z.value};
        double m[3];
        to_local(p, raw, m);
        const double d[3] = {lm.x.is_uncertain ? 1.0 / std::fabs(lm.x.stddev) : 1e6, lm.y.is_uncertain ? 1.0 / std::fabs(lm.y.stddev) : 1e6,
                             lm.z.is_uncertain ? 1.0 / std::fabs(lm.z.stddev) : 1e6};
        put(SBA_PRIOR_POINT, (int32_t)e, d, m);
    }
    for (size_t e = 0; e < p.kept_images.size(); e++) {   // imagealignementsbamodule.cpp:240-286
        const ImageBlock& im = p.images[p.kept_images[e]];
        if (im.fixed) continue;
        if (im.x.is_uncertain && im.y.is_uncertain && im.z.is_uncertain) {
            const double d[3] = {1.0 / std::fabs(im.x.stddev), 1.0 / std::fabs(im.y.stddev), 1.0 / std::fabs(im.z.stddev)};
            const double m[3] = {im.x.value, im.y.value, im.z.value};
            put(SBA_PRIOR_CAM_T, (int32_t)e, d, m);
        }
        if (im.rx.is_uncertain && im.ry.is_uncertain && im.rz.is_uncertain) {
            const double d[3] = {1.0 / std::fabs(im.rx.stddev), 1.0 / std::fabs(im.ry.stddev), 1.0 / std::fabs(im.rz.stddev)};
            const double m[3] = {im.rx.value, im.ry.value, im.rz.value};
            put(SBA_PRIOR_CAM_R, (int32_t)e, d, m);
        }
    }
    return n;
}

// intrinsic blocks of the kept cameras: values, constant mask, priors (pinholecameraprojectormodule.cpp:302-533)
int32_t emit_intrinsics(const sba_project& p, int32_t fixed_flags, double* intr10, uint8_t* mask, int32_t* pm, int32_t* pb, double* A9, double* b3) {
    int32_t n = 0;
    auto prior = [&](int32_t model, int32_t block, int dim, const double* sigma, const double* mean) {
        if (pm) {
            pm[n] = model; pb[n] = block;
            for (int i = 0; i < 9; i++) A9[9 * (size_t)n + i] = 0.0;
            for (int i = 0; i < 3; i++) b3[3 * (size_t)n + i] = 0.0;
            // "float sigma = ...stddev(); A(i,i) = 1/sigma;" -- the division happens in float there
            for (int i = 0; i < dim; i++) { A9[9 * (size_t)n + (size_t)i * dim + i] = (double)(1.0f / (float)sigma[i]); b3[3 * (size_t)n + i] = mean[i]; }
        }
        n++;
    };
    for (size_t m = 0; m < p.kept_cameras.size(); m++) {
        const CameraBlock& c = p.cameras[p.kept_cameras[m]];
        const bool all_const = c.fixed || (fixed_flags & 1);   // FixedParameter::CameraInternal
        double v[10];
        uint8_t mk = 0;
        const bool opp = c.ocx.is_set && c.ocy.is_set;
        v[1] = opp ? c.ocx.value : c.cx.value; v[2] = opp ? c.ocy.value : c.cy.value;
        v[0] = c.of.is_set ? c.of.value : c.f.value;
        if (all_const) mk |= 1 | 2;
        else {
            // the reference adds the pp prior before it looks at the constant flags (:327-345); on a constant block it only
            // shifts the cost by a constant, so it is emitted for free blocks only
            if (c.cx.is_set && c.cy.is_set && c.cx.is_uncertain && c.cy.is_uncertain) {
                const double s[2] = {(double)(float)c.cx.stddev, (double)(float)c.cy.stddev}, b[2] = {c.cx.value, c.cy.value};
                prior((int32_t)m, 1, 2, s, b);
            }
            if (c.f.is_set && c.f.is_uncertain) {
                const double s[1] = {(double)(float)c.f.stddev}, b[1] = {c.f.value};
                prior((int32_t)m, 0, 1, s, b);
            }
        }
        const bool ok = c.ok1.is_set && c.ok2.is_set && c.ok3.is_set;
        v[3] = ok ? c.ok1.value : c.k1.value; v[4] = ok ? c.ok2.value : c.k2.value; v[5] = ok ? c.ok3.value : c.k3.value;
        if (!c.use_rd) { v[3] = v[4] = v[5] = 0.0; mk |= 4; }
        else if (all_const) mk |= 4;
        else if (c.k1.is_set && c.k2.is_set && c.k3.is_set && c.k1.is_uncertain && c.k2.is_uncertain && c.k3.is_uncertain) {
            const double s[3] = {(double)(float)c.k1.stddev, (double)(float)c.k2.stddev, (double)(float)c.k3.stddev}, b[3] = {c.k1.value, c.k2.value, c.k3.value};
            prior((int32_t)m, 2, 3, s, b);
        }
        const bool ot = c.op1.is_set && c.op2.is_set;
        v[6] = ot ? c.op1.value : c.p1.value; v[7] = ot ? c.op2.value : c.p2.value;
        if (!c.use_td) { v[6] = v[7] = 0.0; mk |= 8; }
        else if (all_const) mk |= 8;
        else if (c.p1.is_set && c.p2.is_set && c.p1.is_uncertain && c.p2.is_uncertain) {
            const double s[2] = {(double)(float)c.p1.stddev, (double)(float)c.p2.stddev}, b[2] = {c.p1.value, c.p2.value};
            prior((int32_t)m, 3, 2, s, b);
        }
        const bool oB = c.oB1.is_set && c.oB2.is_set;
        v[8] = oB ? c.oB1.value : c.B1.value; v[9] = oB ? c.oB2.value : c.B2.value;
        if (!c.use_sd) { v[8] = v[9] = 0.0; mk |= 16; }
        else if (all_const) mk |= 16;
        else if (c.B1.is_set && c.B2.is_set && c.B1.is_uncertain && c.B2.is_uncertain) {
            const double s[2] = {(double)(float)c.B1.stddev, (double)(float)c.B2.stddev}, b[2] = {c.B1.value, c.B2.value};
            prior((int32_t)m, 4, 2, s, b);
        }
        if (intr10) for (int i = 0; i < 10; i++) intr10[10 * m + i] = v[i];
        if (mask) mask[m] = mk;
    }
    return n;
}

}  // namespace

extern "C" {

const char* sba_project_error(void) { return g_proj_error.c_str(); }

int sba_project_open(const char* path, sba_project** out) {
    if (!path || !out) return proj_fail("sba_project_open: null argument");
    *out = nullptr;
    FILE* f = std::fopen(path, "rb");
    if (!f) return proj_fail(std::string("cannot open ") + path);
    std::vector<char> buf;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n < 0) { std::fclose(f); return proj_fail(std::string("cannot read ") + path); }
    buf.resize((size_t)n + 1);
    const size_t got = std::fread(buf.data(), 1, (size_t)n, f);
    std::fclose(f);
    buf[got] = 0;
    sba_project* p = new (std::nothrow) sba_project();
    if (!p) return proj_fail("out of memory");
    Parser ps{buf.data(), buf.data() + got, {}, 0};
    if (!ps.value(p->doc)) {
        const std::string e = ps.err;
        const long at = (long)(ps.p - buf.data());
        delete p;
        return proj_fail(std::string(path) + ": JSON error at byte " + std::to_string(at) + ": " + e);
    }
    ps.ws();
    if (ps.p != ps.end) { delete p; return proj_fail(std::string(path) + ": trailing characters after the JSON document"); }
    if (p->doc.type != JVal::Obj) { delete p; return proj_fail(std::string(path) + ": the document is not a JSON object (Project::load rejects it too)"); }
    int rc = parse_blocks(*p);
    if (rc == SBA_OK) rc = build_problem(*p);
    if (rc != SBA_OK) { delete p; return rc; }
    p->n_priors = emit_priors(*p, nullptr, nullptr, nullptr, nullptr);
    p->n_intr_priors = emit_intrinsics(*p, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
    *out = p;
    return SBA_OK;
}

void sba_project_close(sba_project* p) { delete p; }

int sba_project_sizes(const sba_project* p, int32_t* n_models, int32_t* n_images, int32_t* n_landmarks, int32_t* n_obs,
                      int32_t* n_priors, int32_t* n_intrinsics_priors, int32_t* n_images_without_camera) {
    if (!p) return proj_fail("sba_project_sizes: null handle");
    if (n_models) *n_models = (int32_t)p->kept_cameras.size();
    if (n_images) *n_images = (int32_t)p->kept_images.size();
    if (n_landmarks) *n_landmarks = (int32_t)p->kept_landmarks.size();
    if (n_obs) *n_obs = (int32_t)p->obs_cam.size();
    if (n_priors) *n_priors = p->n_priors;
    if (n_intrinsics_priors) *n_intrinsics_priors = p->n_intr_priors;   // upper bound: with SBA_PROJECT_FIX_CAMERA_INTERNAL none are emitted
    if (n_images_without_camera) *n_images_without_camera = p->skipped_images_without_camera;
    return SBA_OK;
}

int sba_project_ids(const sba_project* p, int64_t* model_id, int64_t* image_id, int64_t* landmark_id) {
    if (!p) return proj_fail("sba_project_ids: null handle");
    if (model_id) for (size_t i = 0; i < p->kept_cameras.size(); i++) model_id[i] = p->cameras[p->kept_cameras[i]].id;
    if (image_id) for (size_t i = 0; i < p->kept_images.size(); i++) image_id[i] = p->images[p->kept_images[i]].id;
    if (landmark_id) for (size_t i = 0; i < p->kept_landmarks.size(); i++) landmark_id[i] = p->landmarks[p->kept_landmarks[i]].id;
    return SBA_OK;
}

int sba_project_convert(const sba_project* p, int32_t fixed_flags, double* intr10, uint8_t* model_fixed_mask, double* raxis, double* t,
                        int32_t* cam_model, uint8_t* cam_fixed, double* pts, uint8_t* pt_fixed, int32_t* obs_cam, int32_t* obs_pt,
                        double* obs_uv, double* obs_info, int32_t* prior_kind, int32_t* prior_index, double* prior_A9, double* prior_b3,
                        int32_t* ipr_model, int32_t* ipr_block, double* ipr_A9, double* ipr_b3, int32_t* n_intrinsics_priors) {
    if (!p) return proj_fail("sba_project_convert: null handle");
    if ((prior_kind || prior_index || prior_A9 || prior_b3) && !(prior_kind && prior_index && prior_A9 && prior_b3))
        return proj_fail("sba_project_convert: the four prior arrays go together");
    if ((ipr_model || ipr_block || ipr_A9 || ipr_b3) && !(ipr_model && ipr_block && ipr_A9 && ipr_b3))
        return proj_fail("sba_project_convert: the four intrinsics-prior arrays go together");
    const int32_t nip = emit_intrinsics(*p, fixed_flags, intr10, model_fixed_mask, ipr_model, ipr_block, ipr_A9, ipr_b3);
    if (n_intrinsics_priors) *n_intrinsics_priors = nip;
    std::vector<int32_t> model_of(p->cameras.size(), -1);
    for (size_t m = 0; m < p->kept_cameras.size(); m++) model_of[p->kept_cameras[m]] = (int32_t)m;
    std::map<long long, int32_t> cam_of;
    for (size_t i = 0; i < p->cameras.size(); i++) cam_of.emplace(p->cameras[i].id, (int32_t)i);
    for (size_t e = 0; e < p->kept_images.size(); e++) {   // modularsbasolver.cpp:265-302
        const ImageBlock& im = p->images[p->kept_images[e]];
        const bool pr = im.rx.is_set && im.ry.is_set && im.rz.is_set, pt = im.x.is_set && im.y.is_set && im.z.is_set;
        // without a prior the reference copies an uninitialised std::array here; zero is what this reader uses
        const double rp[3] = {pr ? im.rx.value : 0.0, pr ? im.ry.value : 0.0, pr ? im.rz.value : 0.0};
        const double tp[3] = {pt ? im.x.value : 0.0, pt ? im.y.value : 0.0, pt ? im.z.value : 0.0};
        const bool use_prior_r = im.fixed || !im.orot.is_set, use_prior_t = im.fixed || !im.op.is_set;
        for (int i = 0; i < 3; i++) {
            if (raxis) raxis[3 * e + i] = use_prior_r ? rp[i] : im.orot.v[i];
            if (t) t[3 * e + i] = use_prior_t ? tp[i] : im.op.v[i];
        }
        if (cam_model) cam_model[e] = model_of[cam_of.find(im.camera)->second];
        if (cam_fixed) cam_fixed[e] = im.fixed ? 1 : 0;
    }
    for (size_t e = 0; e < p->kept_landmarks.size(); e++) {   // landmarkssbamodule.cpp:143-171
        const LandmarkBlock& lm = p->landmarks[p->kept_landmarks[e]];
        double x[3] = {0.0, 0.0, 0.0};
        if (lm.op.is_set) to_local(*p, lm.op.v, x);
        else if (lm.x.is_set && lm.y.is_set && lm.z.is_set) { const double raw[3] = {lm.x.value, lm.y.value, lm.z.value}; to_local(*p, raw, x); }
        if (pts) for (int i = 0; i < 3; i++) pts[3 * e + i] = x[i];
        if (pt_fixed) pt_fixed[e] = lm.fixed ? 1 : 0;
    }
    const size_t no = p->obs_cam.size();
    if (obs_cam) std::copy(p->obs_cam.begin(), p->obs_cam.end(), obs_cam);
    if (obs_pt) std::copy(p->obs_pt.begin(), p->obs_pt.end(), obs_pt);
    if (obs_uv) std::copy(p->obs_uv.begin(), p->obs_uv.begin() + 2 * no, obs_uv);
    if (obs_info) std::copy(p->obs_info.begin(), p->obs_info.begin() + 4 * no, obs_info);
    if (prior_kind) emit_priors(*p, prior_kind, prior_index, prior_A9, prior_b3);
    return SBA_OK;
}

int sba_project_write_results(sba_project* p, const double* raxis, const double* t, const double* pts, const double* intr10) {
    if (!p) return proj_fail("sba_project_write_results: null handle");
    auto f32 = [](double v) { return (double)(float)v; };   // the modules store static_cast<float>(value)
    if (raxis && t)
        if (JVal* arr = p->doc.get(kImageClass))
            for (size_t e = 0; e < p->kept_images.size(); e++) {   // imagealignementsbamodule.cpp:446-490
                ImageBlock& im = p->images[p->kept_images[e]];
                if (im.fixed) continue;
                for (int i = 0; i < 3; i++) { im.op.v[i] = f32(t[3 * e + i]); im.orot.v[i] = f32(raxis[3 * e + i]); }
                im.op.is_set = im.orot.is_set = true;
                JVal& o = arr->arr[im.json_index];
                o.set("op", float_group3_json(im.op.v));
                o.set("or", float_group3_json(im.orot.v));
            }
    if (pts)
        if (JVal* arr = p->doc.get(kLandmarkClass))
            for (size_t e = 0; e < p->kept_landmarks.size(); e++) {   // landmarkssbamodule.cpp:207-262 (no CRS: the local-frame value is stored as it is)
                LandmarkBlock& lm = p->landmarks[p->kept_landmarks[e]];
                for (int i = 0; i < 3; i++) lm.op.v[i] = f32(pts[3 * e + i]);
                lm.op.is_set = true;
                arr->arr[lm.json_index].set("op", float_group3_json(lm.op.v));
            }
    if (intr10)
        if (JVal* arr = p->doc.get(kCameraClass))
            for (size_t m = 0; m < p->kept_cameras.size(); m++) {   // pinholecameraprojectormodule.cpp:535-574
                CameraBlock& c = p->cameras[p->kept_cameras[m]];
                if (c.fixed) continue;
                JVal& o = arr->arr[c.json_index];
                const double* v = intr10 + 10 * m;
                static const char* const okeys[] = {"of", "opix_ratio", "ocx", "ocy", "ok1", "ok2", "ok3", "ok4", "ok5", "ok6", "op1", "op2", "oB1", "oB2"};
                for (const char* k : okeys) o.set(k, float_param_unset_json());   // Camera::clearOptimized
                c.of = c.ocx = c.ocy = c.ok1 = c.ok2 = c.ok3 = c.op1 = c.op2 = c.oB1 = c.oB2 = FloatParam();
                auto put = [&](const char* key, FloatParam& fp, double val) {
                    fp = FloatParam(); fp.is_set = true; fp.value = f32(val);
                    o.set(key, float_param_json(fp.value));
                };
                put("of", c.of, v[0]); put("ocx", c.ocx, v[1]); put("ocy", c.ocy, v[2]);
                if (c.use_rd) { put("ok1", c.ok1, v[3]); put("ok2", c.ok2, v[4]); put("ok3", c.ok3, v[5]); }
                if (c.use_td) { put("op1", c.op1, v[6]); put("op2", c.op2, v[7]); }
                if (c.use_sd) { put("oB1", c.oB1, v[8]); put("oB2", c.oB2, v[9]); }
            }
    return SBA_OK;
}

int sba_project_save(const sba_project* p, const char* path) {
    if (!p || !path) return proj_fail("sba_project_save: null argument");
    std::string out;
    write_value(out, p->doc, 0);
    out += '\n';
    FILE* f = std::fopen(path, "wb");
    if (!f) return proj_fail(std::string("cannot write ") + path);
    const size_t w = std::fwrite(out.data(), 1, out.size(), f);
    const int c = std::fclose(f);
    if (w != out.size() || c != 0) return proj_fail(std::string("short write to ") + path);
    return SBA_OK;
}

}  // extern "C"
