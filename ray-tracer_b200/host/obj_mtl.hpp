// obj_mtl.hpp -- SURVEY.md 8(f) rows 1 and 3: OBJ + MTL loading with per-face materials and `vt`
// texture coordinates, and the parsed-texture file reader (text format of the reference's
// textures/parse_textures.py plus a binary fast path).
//
// The reference's loader reads `v` and `f` lines only (src/obj_read.cu:94-146), gives a whole mesh
// ONE material and ignores .mtl files; meshes get the texture points of the Vec3 constructors
// (src/main.cu:127-148).  This header adds what BASELINE config 3 asks for, written -- like
// scenes.hpp, which includes it -- ONLY against the reference's public names
//   read_file, split_string, Matrix/EnlargementMatrix/RotationMatrix, Vec3, Vec2, Vertex,
//   Triangle(Vertex, Vertex, Vertex, Material), Texture::create_*, Material, Object::create_mesh,
//   ReadOnlyDeviceArray<Triangle>
// so the same file compiles against rtb_compat.hpp (product), against the reference's sources for
// the host (oracle/ref_host.cpp) and against its CUDA sources (baseline/ref_cuda.cu): the scene the
// loader produces is rendered by the reference's own kernel in the parity tests.
//
// Mapping (documented in DESIGN.md 4.5):
//   Ke != 0            -> EMISSIVE, colour Ke / max(Ke), strength max(Ke)
//   else d < 1         -> REFRACTIVE, refractive_index Ni, texture from Kd
//   else               -> STANDARD, smoothness sqrt(Ns / 1000) clamped to [0, 1] (Blender writes
//                         Ns = 1000 (1 - roughness)^2), texture map_Kd (IMAGE, looked up by file
//                         name in the parsed-texture file) or Kd (COLOUR)
//   usemtl groups      -> one MESH object per material, in order of first use (creation order is
//                         the tie rule, src/raytracer.cu:36); faces before any usemtl use the
//                         caller's default material
//   f a/b/c            -> vertex a, texture point b (`vt`, Triangle::texture_points); `vn` is parsed
//                         and kept but has no consumer: the path shades with the face normal
//                         (src/objects.cu:175-186)
//   4-vertex faces     -> triangles (1,2,3) and (1,4,3), the split of Quad (src/objects.cu:244-253)
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace rtb_scenes {

// where this repo's own fixtures live (ray-tracer_b200/assets): $RTB_ASSETS_DIR, else relative to cwd
inline std::string &assets_dir() {
    static std::string dir = [] {
        const char *e = std::getenv("RTB_ASSETS_DIR");
        return std::string(e && *e ? e : "ray-tracer_b200/assets");
    }();
    return dir;
}

// split_string gives every token after the first a leading space (src/obj_read.cu:35) and keeps the
// empty tokens between repeated separators: trim, and drop the empty ones and a trailing '\r'
inline std::string trimmed(const std::string &t) {
    size_t a = 0, b = t.size();
    while (a < b && (t[a] == ' ' || t[a] == '\t')) a++;
    while (b > a && (t[b - 1] == ' ' || t[b - 1] == '\t' || t[b - 1] == '\r' || t[b - 1] == '\n')) b--;
    return t.substr(a, b - a);
}

inline std::vector<std::string> words(const std::string &line) {
    std::vector<std::string> out;
    for (std::string &t : split_string(line, ' ')) {
        std::string w = trimmed(t);
        if (!w.empty()) out.push_back(w);
    }
    return out;
}

// ---------------------------------------------------------------------------------------------
// parsed-texture files
// ---------------------------------------------------------------------------------------------
struct ParsedImage {
    int width = 0, height = 0;
    std::vector<Vec3> rgb;
};

// Text format == textures/parse_textures.py:29-58 (count line; per image: name, width, height,
// one line of `x / 256` values each followed by a space), read like ImageTexture::parse_file
// (src/main.cu:58-90).
inline bool load_parsed_texture_text(const std::string &file, const std::string &name, ParsedImage &out) {
    FILE *probe = std::fopen(file.c_str(), "rb");
    if (!probe) return false;
    std::fclose(probe);
    std::vector<std::string> lines = read_file(file);
    for (size_t i = 0; i + 3 < lines.size(); i++) {
        if (trimmed(lines[i]) != name) continue;
        out.width = std::stoi(lines[i + 1]);
        out.height = std::stoi(lines[i + 2]);
        std::vector<std::string> tok = split_string(lines[i + 3], ' ');
        out.rgb.clear();
        for (size_t k = 0; k + 2 < tok.size(); k += 3)  // the line ends with a space: one blank token follows the last value
            out.rgb.push_back(Vec3(std::stof(tok[k]), std::stof(tok[k + 1]), std::stof(tok[k + 2])));
        return true;
    }
    return false;
}

// Binary fast path (ray-tracer_b200/textures.py writes it): little endian
//   char magic[8] = "RTBTEX1\0"; u32 count; per image { u32 name_len; name; u32 width; u32 height;
//   f32 rgb[width * height * 3] }  -- the same float values the text format holds (x / 256 is exact).
inline bool load_parsed_texture_binary(const std::string &file, const std::string &name, ParsedImage &out) {
    FILE *f = std::fopen(file.c_str(), "rb");
    if (!f) return false;
    bool found = false;
    char magic[8];
    uint32_t count = 0;
    if (std::fread(magic, 1, 8, f) == 8 && std::memcmp(magic, "RTBTEX1\0", 8) == 0 && std::fread(&count, 4, 1, f) == 1) {
        for (uint32_t i = 0; i < count && !found; i++) {
            uint32_t len = 0, w = 0, h = 0;
            if (std::fread(&len, 4, 1, f) != 1 || len > 4096) break;
            std::string nm(len, '\0');
            if (len && std::fread(&nm[0], 1, len, f) != len) break;
            if (std::fread(&w, 4, 1, f) != 1 || std::fread(&h, 4, 1, f) != 1) break;
            size_t n = (size_t)w * h * 3;
            if (nm != name) {
                if (std::fseek(f, (long)(n * 4), SEEK_CUR) != 0) break;
                continue;
            }
            std::vector<float> v(n);
            if (n && std::fread(v.data(), 4, n, f) != n) break;
            out.width = (int)w;
            out.height = (int)h;
            out.rgb.clear();
            out.rgb.reserve((size_t)w * h);
            for (size_t k = 0; k < n; k += 3) out.rgb.push_back(Vec3(v[k], v[k + 1], v[k + 2]));
            found = true;
        }
    }
    std::fclose(f);
    return found;
}

// <dir>/parsed_textures.bin if it holds the image, else <dir>/parsed_textures.txt
inline ParsedImage load_parsed_texture(const std::string &dir, const std::string &name) {
    ParsedImage img;
    if (load_parsed_texture_binary(dir + "/parsed_textures.bin", name, img)) return img;
    if (load_parsed_texture_text(dir + "/parsed_textures.txt", name, img)) return img;
    throw std::runtime_error("Image file not found.\n");  // the reference's message, src/main.cu:72
}

// ---------------------------------------------------------------------------------------------
// .mtl
// ---------------------------------------------------------------------------------------------
struct MtlEntry {
    std::string name;
    float Kd[3] = {0.8f, 0.8f, 0.8f};
    float Ke[3] = {0, 0, 0};
    float Ns = 0, Ni = 1, d = 1;
    std::string map_Kd;
};

class MtlLibrary {
  public:
    std::vector<MtlEntry> entries;

    MtlLibrary() {}
    explicit MtlLibrary(const std::string &filename) {
        for (const std::string &line : read_file(filename)) {
            std::vector<std::string> t = words(line);
            if (t.empty() || t[0][0] == '#') continue;
            if (t[0] == "newmtl" && t.size() > 1) {
                entries.push_back(MtlEntry());
                entries.back().name = t[1];
            } else if (!entries.empty()) {
                MtlEntry &e = entries.back();
                if (t[0] == "Kd" && t.size() > 3) { for (int k = 0; k < 3; k++) e.Kd[k] = std::stof(t[1 + k]); }
                else if (t[0] == "Ke" && t.size() > 3) { for (int k = 0; k < 3; k++) e.Ke[k] = std::stof(t[1 + k]); }
                else if (t[0] == "Ns" && t.size() > 1) e.Ns = std::stof(t[1]);
                else if (t[0] == "Ni" && t.size() > 1) e.Ni = std::stof(t[1]);
                else if (t[0] == "d" && t.size() > 1) e.d = std::stof(t[1]);
                else if (t[0] == "Tr" && t.size() > 1) e.d = 1.0f - std::stof(t[1]);
                else if (t[0] == "map_Kd" && t.size() > 1) e.map_Kd = t.back();  // options, if any, precede the file name
            }
        }
    }

    const MtlEntry *find(const std::string &name) const {
        for (const MtlEntry &e : entries)
            if (e.name == name) return &e;
        return nullptr;
    }

    // texture_dir holds parsed_textures.{bin,txt} (map_Kd images are looked up there by file name)
    static Material to_material(const MtlEntry &e, const std::string &texture_dir) {
        float ke_max = std::fmax(e.Ke[0], std::fmax(e.Ke[1], e.Ke[2]));
        if (ke_max > 0) return make_emissive(Vec3(e.Ke[0] / ke_max, e.Ke[1] / ke_max, e.Ke[2] / ke_max), ke_max);
        Texture tex = Texture::create_const_colour(Vec3(e.Kd[0], e.Kd[1], e.Kd[2]));
        if (e.d < 1) return make_refractive(tex, e.Ni);
        if (!e.map_Kd.empty()) {
            std::string file = e.map_Kd;
            size_t slash = file.find_last_of("/\\");
            if (slash != std::string::npos) file = file.substr(slash + 1);
            ParsedImage img = load_parsed_texture(texture_dir, file);
            tex = Texture::create_image(img.width, img.height, img.rgb);
        }
        float smooth = std::sqrt(std::fmax(e.Ns, 0.0f) / 1000.0f);
        return make_standard(tex, smooth > 1 ? 1.0f : smooth);
    }
};

// ---------------------------------------------------------------------------------------------
// .obj with vt / vn / usemtl / mtllib
// ---------------------------------------------------------------------------------------------
class ObjMtlMesh {
  public:
    Matrix vertex_mat;                         // 3 x num_vertices, transformed like ObjFileMesh::vertex_mat
    std::vector<Vec2> texture_points;          // vt
    std::vector<Vec3> normals;                 // vn (kept, unused by the path)
    std::vector<std::string> material_names;   // usemtl names in order of first use ("" = faces before any usemtl)
    MtlLibrary library;

    struct Corner { int v, vt, vn; };
    struct Face { std::vector<Corner> corners; int group; };
    std::vector<Face> faces;

    explicit ObjMtlMesh(const std::string &filename) {
        std::string dir = ".";
        size_t slash = filename.find_last_of("/\\");
        if (slash != std::string::npos) dir = filename.substr(0, slash);
        std::vector<std::vector<float>> vertices;
        int group = -1;
        for (const std::string &line : read_file(filename)) {
            std::vector<std::string> t = words(line);
            if (t.empty() || t[0][0] == '#') continue;
            if (t[0] == "v" && t.size() > 3) {
                vertices.push_back({std::stof(t[1]), std::stof(t[2]), std::stof(t[3])});
            } else if (t[0] == "vt" && t.size() > 2) {
                texture_points.push_back(Vec2(std::stof(t[1]), std::stof(t[2])));
            } else if (t[0] == "vn" && t.size() > 3) {
                normals.push_back(Vec3(std::stof(t[1]), std::stof(t[2]), std::stof(t[3])));
            } else if (t[0] == "mtllib" && t.size() > 1) {
                // a library that is not there is not an error: its materials fall back to the caller's default
                FILE *probe = std::fopen((dir + "/" + t[1]).c_str(), "rb");
                if (probe) {
                    std::fclose(probe);
                    library = MtlLibrary(dir + "/" + t[1]);
                }
            } else if (t[0] == "usemtl" && t.size() > 1) {
                group = group_of(t[1]);
            } else if (t[0] == "f") {
                if (group < 0) group = group_of("");
                Face face;
                face.group = group;
                for (size_t i = 1; i < t.size(); i++) {
                    std::vector<std::string> idx = split_string(t[i], '/');
                    Corner c{resolve(idx, 0, (int)vertices.size()), resolve(idx, 1, (int)texture_points.size()),
                             resolve(idx, 2, (int)normals.size())};
                    if (c.v < 0) throw std::runtime_error("OBJ face refers to a vertex that does not exist.\n");
                    face.corners.push_back(c);
                }
                if (face.corners.size() != 3 && face.corners.size() != 4)
                    throw std::logic_error("Only triangle or quad meshes are supported.\n");  // src/main.cu:141
                faces.push_back(face);
            }
        }
        vertex_mat = Matrix(vertices).transpose();
    }

    // the three transforms of ObjFileMesh (src/obj_read.cu:58-85), same matrix arithmetic
    void enlarge(float scale_fact) { vertex_mat = EnlargementMatrix(scale_fact, 3) * vertex_mat; }
    void rotate(float x_angle, float y_angle, float z_angle) {
        RotationMatrix x_rot(RotationMatrix::X_AXIS, x_angle);
        RotationMatrix y_rot(RotationMatrix::Y_AXIS, y_angle);
        RotationMatrix z_rot(RotationMatrix::Z_AXIS, z_angle);
        vertex_mat = x_rot * y_rot * z_rot * vertex_mat;
    }
    void translate(float offset_x, float offset_y, float offset_z) {
        for (int i = 0; i < vertex_mat.cols; i++) {
            vertex_mat.items[0][i] += offset_x;
            vertex_mat.items[1][i] += offset_y;
            vertex_mat.items[2][i] += offset_z;
        }
    }

    // One MESH object per material group, in order of first use.  Faces of a group keep file order.
    std::vector<Object> create_objects(Material default_material, const std::string &texture_dir) {
        std::vector<Object> objects;
        for (size_t g = 0; g < material_names.size(); g++) {
            const MtlEntry *e = library.find(material_names[g]);
            Material mat = e ? MtlLibrary::to_material(*e, texture_dir) : default_material;
            std::vector<Triangle> triangles;
            for (const Face &f : faces) {
                if (f.group != (int)g) continue;
                Vertex v1 = vertex(f.corners[0]), v2 = vertex(f.corners[1]), v3 = vertex(f.corners[2]);
                triangles.push_back(Triangle(v1, v2, v3, mat));
                if (f.corners.size() == 4) triangles.push_back(Triangle(v1, vertex(f.corners[3]), v3, mat));
            }
            if (triangles.empty()) continue;
            ReadOnlyDeviceArray<Triangle> device_array(triangles);
            objects.push_back(Object::create_mesh(triangles, device_array.device_pointer, mat));
        }
        return objects;
    }

  private:
    int group_of(const std::string &name) {
        for (size_t i = 0; i < material_names.size(); i++)
            if (material_names[i] == name) return (int)i;
        material_names.push_back(name);
        return (int)material_names.size() - 1;
    }

    // OBJ indices are 1-based; negative ones count back from the end; a missing field gives -1
    static int resolve(const std::vector<std::string> &idx, size_t field, int count) {
        if (field >= idx.size()) return -1;
        std::string s = trimmed(idx[field]);
        if (s.empty()) return -1;
        int i = std::stoi(s);
        i = i < 0 ? count + i : i - 1;
        return (i >= 0 && i < count) ? i : -1;
    }

    Vertex vertex(const Corner &c) {
        // aggregate initialisation: the reference's Vec2() is device-only (src/utils.cu:171)
        return Vertex{Vec3(vertex_mat.items[0][c.v], vertex_mat.items[1][c.v], vertex_mat.items[2][c.v]),
                      c.vt >= 0 ? texture_points[c.vt] : Vec2(0, 0)};
    }
};

}  // namespace rtb_scenes
