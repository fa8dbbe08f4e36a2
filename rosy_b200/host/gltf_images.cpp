// gltf_images.cpp - see gltf_images.h.  A small recursive-descent JSON reader (RFC 8259) stands in
// for fastgltf: only "images", "textures" and "materials" are looked at.
#include "gltf_images.h"

#include <algorithm>
#include <cstdlib>
#include <format>
#include <fstream>
#include <map>
#include <sstream>

namespace {

struct jvalue {
    enum kind_t { null_k, bool_k, number_k, string_k, array_k, object_k } kind{null_k};
    bool b{false};
    double num{0.0};
    std::string str;
    std::vector<jvalue> arr;
    std::vector<std::pair<std::string, jvalue>> obj;

    const jvalue* get(const char* key) const {
        if (kind != object_k) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    // non-negative integer member, or -1
    long long index_of(const char* key) const {
        const jvalue* v = get(key);
        if (!v || v->kind != number_k || v->num < 0 || v->num != (double)(long long)v->num) return -1;
        return (long long)v->num;
    }
};

class jparser {
public:
    explicit jparser(const std::string& text) : s_(text) {}
    bool parse(jvalue& out, std::string& err) {
        try {
            ws();
            value(out, 0);
            ws();
            if (p_ != s_.size()) fail("trailing characters");
            return true;
        } catch (const std::string& e) {
            err = std::format("{} at byte {}", e, p_);
            return false;
        }
    }

private:
    const std::string& s_;
    size_t p_{0};
    [[noreturn]] void fail(const char* what) { throw std::string(what); }
    void ws() {
        while (p_ < s_.size() && (s_[p_] == ' ' || s_[p_] == '\t' || s_[p_] == '\n' || s_[p_] == '\r')) p_++;
    }
    bool lit(const char* w) {
        const size_t n = std::char_traits<char>::length(w);
        if (s_.compare(p_, n, w) != 0) return false;
        p_ += n;
        return true;
    }
    static void utf8(std::string& o, unsigned cp) {
        if (cp < 0x80) {
            o += (char)cp;
        } else if (cp < 0x800) {
            o += (char)(0xC0 | cp >> 6);
            o += (char)(0x80 | (cp & 63));
        } else if (cp < 0x10000) {
            o += (char)(0xE0 | cp >> 12);
            o += (char)(0x80 | ((cp >> 6) & 63));
            o += (char)(0x80 | (cp & 63));
        } else {
            o += (char)(0xF0 | cp >> 18);
            o += (char)(0x80 | ((cp >> 12) & 63));
            o += (char)(0x80 | ((cp >> 6) & 63));
            o += (char)(0x80 | (cp & 63));
        }
    }
    unsigned hex4() {
        if (p_ + 4 > s_.size()) fail("short \\u escape");
        unsigned v = 0;
        for (int i = 0; i < 4; i++) {
            const char c = s_[p_++];
            v <<= 4;
            if (c >= '0' && c <= '9') v |= (unsigned)(c - '0');
            else if (c >= 'a' && c <= 'f') v |= (unsigned)(c - 'a' + 10);
            else if (c >= 'A' && c <= 'F') v |= (unsigned)(c - 'A' + 10);
            else fail("bad \\u escape");
        }
        return v;
    }
    void string(std::string& o) {
        if (s_[p_] != '"') fail("expected string");
        p_++;
        for (;;) {
            if (p_ >= s_.size()) fail("unterminated string");
            const char c = s_[p_++];
            if (c == '"') return;
            if (c != '\\') {
                o += c;
                continue;
            }
            if (p_ >= s_.size()) fail("unterminated escape");
            const char e = s_[p_++];
            switch (e) {
                case '"': o += '"'; break;
                case '\\': o += '\\'; break;
                case '/': o += '/'; break;
                case 'b': o += '\b'; break;
                case 'f': o += '\f'; break;
                case 'n': o += '\n'; break;
                case 'r': o += '\r'; break;
                case 't': o += '\t'; break;
                case 'u': {
                    unsigned cp = hex4();
                    if (cp >= 0xD800 && cp < 0xDC00 && s_.compare(p_, 2, "\\u") == 0) {
                        p_ += 2;
                        const unsigned lo = hex4();
                        cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
                    }
                    utf8(o, cp);
                    break;
                }
                default: fail("bad escape");
            }
        }
    }
    void value(jvalue& v, int depth) {
        if (depth > 64) fail("nesting too deep");
        if (p_ >= s_.size()) fail("unexpected end");
        const char c = s_[p_];
        if (c == '{') {
            v.kind = jvalue::object_k;
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == '}') {
                p_++;
                return;
            }
            for (;;) {
                ws();
                std::string key;
                string(key);
                ws();
                if (p_ >= s_.size() || s_[p_] != ':') fail("expected ':'");
                p_++;
                ws();
                v.obj.emplace_back(std::move(key), jvalue{});
                value(v.obj.back().second, depth + 1);
                ws();
                if (p_ < s_.size() && s_[p_] == ',') {
                    p_++;
                    continue;
                }
                if (p_ < s_.size() && s_[p_] == '}') {
                    p_++;
                    return;
                }
                fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            v.kind = jvalue::array_k;
            p_++;
            ws();
            if (p_ < s_.size() && s_[p_] == ']') {
                p_++;
                return;
            }
            for (;;) {
                ws();
                v.arr.emplace_back();
                value(v.arr.back(), depth + 1);
                ws();
                if (p_ < s_.size() && s_[p_] == ',') {
                    p_++;
                    continue;
                }
                if (p_ < s_.size() && s_[p_] == ']') {
                    p_++;
                    return;
                }
                fail("expected ',' or ']'");
            }
        }
        if (c == '"') {
            v.kind = jvalue::string_k;
            string(v.str);
            return;
        }
        if (lit("true")) {
            v.kind = jvalue::bool_k;
            v.b = true;
            return;
        }
        if (lit("false")) {
            v.kind = jvalue::bool_k;
            return;
        }
        if (lit("null")) return;
        // number
        const char* start = s_.c_str() + p_;
        char* end = nullptr;
        const double d = std::strtod(start, &end);
        if (end == start) fail("unexpected character");
        v.kind = jvalue::number_k;
        v.num = d;
        p_ += (size_t)(end - start);
    }
};

// "%20" -> ' ' (glTF URIs are percent-encoded, RFC 3986)
std::string percent_decode(const std::string& in) {
    std::string o;
    for (size_t i = 0; i < in.size(); i++) {
        if (in[i] == '%' && i + 2 < in.size()) {
            auto hex = [](char c) { return c >= '0' && c <= '9' ? c - '0' : (c >= 'a' && c <= 'f' ? c - 'a' + 10 : (c >= 'A' && c <= 'F' ? c - 'A' + 10 : -1)); };
            const int h = hex(in[i + 1]), lo = hex(in[i + 2]);
            if (h >= 0 && lo >= 0) {
                o += (char)(h * 16 + lo);
                i += 2;
                continue;
            }
        }
        o += in[i];
    }
    return o;
}

}  // namespace

namespace rosy_packager {

namespace {
// image index of materials[m].<slot>.index -> textures[].source, or -1 (Gltf.cpp:122-178)
long long slot_image(const jvalue& textures, const jvalue* slot) {
    if (!slot) return -1;
    const long long t = slot->index_of("index");
    if (t < 0 || textures.kind != jvalue::array_k || (size_t)t >= textures.arr.size()) return -1;
    return textures.arr[(size_t)t].index_of("source");
}
}  // namespace

rosy::result plan_gltf_images(const std::shared_ptr<rosy_logger::log>& l, const std::filesystem::path& gltf_path,
                              gltf_image_plan& plan, bool all_textures) {
    plan = gltf_image_plan{};
    std::ifstream f(gltf_path, std::ios::binary);
    if (!f) {
        l->error(std::format("import - failed to open {}", gltf_path.string()));  // Gltf.cpp:74-78
        return rosy::result::error;
    }
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string text = ss.str();
    jvalue root;
    std::string err;
    if (!jparser(text).parse(root, err) || root.kind != jvalue::object_k) {
        l->error(std::format("import - failed to load {}, {}", gltf_path.string(), err));  // Gltf.cpp:92-94
        return rosy::result::error;
    }
    static const jvalue empty_array = [] { jvalue v; v.kind = jvalue::array_k; return v; }();
    const jvalue* images = root.get("images");
    const jvalue* textures = root.get("textures");
    const jvalue* materials = root.get("materials");
    if (!images || images->kind != jvalue::array_k) images = &empty_array;
    if (!textures || textures->kind != jvalue::array_k) textures = &empty_array;
    if (!materials || materials->kind != jvalue::array_k) materials = &empty_array;

    // IMAGES (Gltf.cpp:97-111): every image must be a URI data source
    std::vector<std::string> uris;
    for (const jvalue& img : images->arr) {
        const jvalue* uri = img.get("uri");
        if (!uri || uri->kind != jvalue::string_k || uri->str.rfind("data:", 0) == 0) {
            l->error("not a URI datasource");
            return rosy::result::error;
        }
        uris.push_back(uri->str);
        std::filesystem::path name{gltf_path};
        name.replace_filename(std::format("{}.dds", uri->str.substr(0, uri->str.find('.'))));
        plan.image_names.push_back(name.string());
    }
    // MATERIALS (Gltf.cpp:113-200): three lists, in material order
    std::vector<uint32_t> lists[3];  // colour, metallic, normal
    for (const jvalue& mat : materials->arr) {
        const jvalue* pbr = mat.get("pbrMetallicRoughness");
        const long long c = pbr ? slot_image(*textures, pbr->get("baseColorTexture")) : -1;
        const long long m = pbr ? slot_image(*textures, pbr->get("metallicRoughnessTexture")) : -1;
        const long long n = slot_image(*textures, mat.get("normalTexture"));
        const long long e = all_textures ? slot_image(*textures, mat.get("emissiveTexture")) : -1;
        const long long o = all_textures ? slot_image(*textures, mat.get("occlusionTexture")) : -1;
        const long long idx[5] = {c, m, n, e, o};
        static const int list_of[5] = {0, 1, 2, 0, 1};  // emissive is colour, occlusion is data
        for (int k = 0; k < 5; k++) {
            if (idx[k] < 0) continue;
            if ((size_t)idx[k] >= uris.size()) {
                l->error(std::format("import - image index {} out of range", idx[k]));
                return rosy::result::error;
            }
            lists[list_of[k]].push_back((uint32_t)idx[k]);
            if (k < 3) plan.reference_calls++;
        }
    }
    // conditioning order (Gltf.cpp:202-286): colour, metallic, normal; each image once per list, and
    // only in the last list that names it
    static const uint8_t type_of[3] = {image_type_color, image_type_metallic_roughness, image_type_normal_map};
    std::map<uint32_t, int> last_list;
    std::map<uint32_t, uint32_t> uses;
    for (int k = 0; k < 3; k++)
        for (uint32_t i : lists[k]) {
            last_list[i] = k;
            uses[i]++;
        }
    uint32_t* counts[3] = {&plan.num_color, &plan.num_metallic, &plan.num_normal};
    for (int k = 0; k < 3; k++) {
        std::vector<uint32_t> v = lists[k];
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        *counts[k] = (uint32_t)v.size();
        for (uint32_t i : v) {
            if (last_list[i] != k) continue;
            gltf_image_job j;
            j.image_index = i;
            j.image_type = type_of[k];
            j.uri = uris[i];
            j.source_path = gltf_path;
            j.source_path.replace_filename(percent_decode(uris[i]));  // Gltf.cpp:222-223
            j.dds_path = j.source_path;
            j.dds_path.replace_extension(".dds");  // Packager.cpp:231
            j.asset_image_name = plan.image_names[i];
            j.uses = uses[i];
            plan.jobs.push_back(j);
        }
    }
    l->info(std::format("num color_images {}", plan.num_color));  // Gltf.cpp:208, :235, :262
    l->info(std::format("num metallic_images {}", plan.num_metallic));
    l->info(std::format("num normal_map_images {}", plan.num_normal));
    return rosy::result::ok;
}

rosy::result condition_gltf_images(const std::shared_ptr<rosy_logger::log>& l, const gltf_image_plan& plan, bool serial) {
    for (const gltf_image_job& j : plan.jobs) {
        rosy::result res;
        if (serial) {
            // the reference's loop bodies, Gltf.cpp:225, :252, :280
            res = j.image_type == image_type_normal_map ? generate_normal_map_texture(l, j.source_path)
                                                        : generate_srgb_texture(l, j.source_path);
        } else {
            res = submit_texture(l, j.source_path, j.image_type);
        }
        if (res != rosy::result::ok) {
            const char* what = j.image_type == image_type_color ? "color" : (j.image_type == image_type_normal_map ? "normal" : "metallic");
            l->info(std::format("error creating {} gltf image: {}", what, static_cast<uint8_t>(res)));  // Gltf.cpp:227, :254, :282
            if (!serial) (void)wait_for_textures(l);  // drain what is already queued
            return res;
        }
    }
    return serial ? rosy::result::ok : wait_for_textures(l);
}

namespace {
std::string json_escape(const std::string& s) {
    std::string o;
    for (const char c : s) {
        if (c == '"' || c == '\\') {
            o += '\\';
            o += c;
        } else if ((unsigned char)c < 0x20) {
            o += std::format("\\u{:04x}", (unsigned)c);
        } else {
            o += c;
        }
    }
    return o;
}
}  // namespace

std::string plan_to_json(const gltf_image_plan& plan) {
    std::string o = "{\n  \"images\": [";
    for (size_t i = 0; i < plan.image_names.size(); i++)
        o += std::format("{}\n    \"{}\"", i ? "," : "", json_escape(plan.image_names[i]));
    o += std::format("\n  ],\n  \"num_color\": {}, \"num_metallic\": {}, \"num_normal\": {}, \"reference_calls\": {},\n  \"jobs\": [",
                     plan.num_color, plan.num_metallic, plan.num_normal, plan.reference_calls);
    for (size_t i = 0; i < plan.jobs.size(); i++) {
        const gltf_image_job& j = plan.jobs[i];
        o += std::format("{}\n    {{\"image_index\": {}, \"image_type\": {}, \"uri\": \"{}\", \"source\": \"{}\", \"dds\": \"{}\", \"name\": \"{}\", \"uses\": {}}}",
                         i ? "," : "", j.image_index, (unsigned)j.image_type, json_escape(j.uri), json_escape(j.source_path.string()),
                         json_escape(j.dds_path.string()), json_escape(j.asset_image_name), j.uses);
    }
    o += "\n  ]\n}\n";
    return o;
}

}  // namespace rosy_packager
