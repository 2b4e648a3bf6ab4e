// .crtscene loader: JSON DOM -> crt::Scene.  Format and validation rules follow the reference loader
// (src/core/crt_json.cpp:541-647 and its helpers); construction rules that decide pixel values are cited inline.
// Deliberate differences (all on inputs where the reference has undefined behaviour):
//   * settings.background_color missing         -> rejected (reference compares against the wrong end iterator, :554)
//   * object.material_index out of range        -> rejected (reference indexes material_triangle_flags unchecked, :210)
//   * triangle vertex index out of range        -> rejected (reference indexes the vertex vector unchecked, crt_mesh.cpp:20)
//   * settings.image_settings not an object     -> rejected (reference asserts, :120)
#include "crt_json.h"

#include <iterator>
#include <string>
#include <unordered_map>

#include "crt_image_stbi.h"
#include "crt_json_reader.h"

namespace crt::json {

namespace {

using jsonreader::Value;

bool read_vector(const Value *v, Vector &out) {
    if (!v || !v->is_array() || v->size() != 3)
        return false;
    if (v->kind != Value::Kind::NumberArray)
        return false; // some element is not a number
    out = Vector{ float(v->numbers[0]), float(v->numbers[1]), float(v->numbers[2]) };
    return true;
}

bool read_matrix(const Value *v, Matrix &out) {
    if (!v || !v->is_array() || v->size() != 9 || v->kind != Value::Kind::NumberArray)
        return false;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            out.data[i][j] = float(v->numbers[size_t(i * 3 + j)]);
    return true;
}

// crt_json.cpp:119-143
std::optional<Camera> read_camera(const Value *camera, const Value *image_settings) {
    if (!camera || !image_settings || !image_settings->is_object())
        return std::nullopt;
    const Value *width = image_settings->find("width"), *height = image_settings->find("height");
    if (!width || !width->is_int() || !height || !height->is_int())
        return std::nullopt;
    if (!camera->is_object())
        return std::nullopt;
    Transform transform;
    if (!read_vector(camera->find("position"), transform.location))
        return std::nullopt;
    if (!read_matrix(camera->find("matrix"), transform.rotation))
        return std::nullopt;
    if (const Value *fov = camera->find("fov_degrees")) {
        if (!fov->is_number())
            return std::nullopt;
        return Camera{ width->as_int(), height->as_int(), fov->as_float(), transform };
    }
    return Camera{ width->as_int(), height->as_int(), transform };
}

struct TextureTable {
    std::vector<Texture> textures;
    std::unordered_map<std::string, size_t> by_name;
};

// crt_json.cpp:375-453.  Any malformed entry voids the WHOLE texture list (the caller falls back to {}).
bool read_textures(const Value &list, const std::filesystem::path &asset_root, TextureTable &out) {
    if (!list.is_array())
        return false;
    if (list.kind == Value::Kind::NumberArray)
        return list.empty(); // non-empty array of numbers: first element is not an object
    for (size_t i = 0; i < list.items.size(); ++i) {
        const Value &v = list.items[i];
        if (!v.is_object())
            return false;
        const Value *name = v.find("name");
        if (!name || !name->is_string())
            return false;
        out.by_name[name->string] = i; // a later duplicate name wins
        const Value *type = v.find("type");
        if (!type || !type->is_string())
            return false;
        Texture tex;
        if (type->string == "albedo") {
            tex.type = TextureType::Albedo;
            if (!read_vector(v.find("albedo"), tex.color_a))
                return false;
        } else if (type->string == "edges") {
            tex.type = TextureType::Edges;
            const Value *w = v.find("edge_width");
            if (!w || !w->is_number())
                return false;
            if (!read_vector(v.find("edge_color"), tex.color_a) || !read_vector(v.find("inner_color"), tex.color_b))
                return false;
            tex.scalar = w->as_float();
        } else if (type->string == "checker") {
            tex.type = TextureType::Checker;
            const Value *a = v.find("color_A"), *b = v.find("color_B"), *s = v.find("square_size");
            if (!a || !b || !s || !s->is_number())
                return false;
            if (!read_vector(a, tex.color_a) || !read_vector(b, tex.color_b))
                return false;
            tex.scalar = s->as_float();
        } else if (type->string == "bitmap") {
            tex.type = TextureType::Bitmap;
            const Value *file = v.find("file_path");
            if (!file || !file->is_string())
                return false;
            // leading '/' is dropped and the rest resolved under asset_root (:358-360)
            std::filesystem::path rel{ std::u8string{ file->string.begin(), file->string.end() } };
            std::optional<Image> image = read_stb(asset_root / rel.relative_path());
            if (!image)
                return false;
            tex.image = std::make_shared<const Image>(std::move(*image));
        } else {
            return false;
        }
        out.textures.push_back(std::move(tex));
    }
    return true;
}

struct MaterialTable {
    std::vector<Material> materials;
    std::vector<TriangleFlags> flags;
};

// crt_json.cpp:460-539
bool read_materials(const Value *list, TextureTable &textures, MaterialTable &out) {
    if (!list || !list->is_array() || list->empty() || list->kind != Value::Kind::Array)
        return false;
    for (const Value &v : list->items) {
        if (!v.is_object())
            return false;
        const Value *type = v.find("type");
        if (!type)
            return false;
        const Value *smooth = v.find("smooth_shading");
        if (!smooth || !smooth->is_bool())
            return false;
        bool cull = false;
        if (const Value *c = v.find("back_face_culling")) {
            if (!c->is_bool())
                return false;
            cull = c->boolean;
        }
        Material m{ MaterialType::Diffuse, -1, 1.0f };
        if (!type->is_string())
            return false;
        if (type->string == "diffuse") m.type = MaterialType::Diffuse;
        else if (type->string == "reflective") m.type = MaterialType::Reflective;
        else if (type->string == "refractive") m.type = MaterialType::Refractive;
        else if (type->string == "constant") m.type = MaterialType::Constant;
        else return false;

        if (m.type != MaterialType::Refractive) {
            const Value *albedo = v.find("albedo");
            if (!albedo)
                return false;
            if (albedo->is_string()) {
                auto it = textures.by_name.find(albedo->string);
                if (it == textures.by_name.end())
                    return false;
                m.albedo_map_texture_index = int(it->second);
            } else {
                // an inline colour becomes an implicit Albedo texture appended to the list (:507-515)
                Texture tex;
                tex.type = TextureType::Albedo;
                if (!read_vector(albedo, tex.color_a))
                    return false;
                m.albedo_map_texture_index = int(textures.textures.size());
                textures.textures.push_back(std::move(tex));
            }
        } else if (const Value *ior = v.find("ior")) { // optional for refractive (:516-525)
            if (!ior->is_number())
                return false;
            m.ior = ior->as_float();
        }
        out.materials.push_back(m);
        out.flags.push_back(TriangleFlags{ uint8_t(smooth->boolean), uint8_t(cull) });
    }
    return true;
}

// crt_json.cpp:150-218 + crt_mesh.cpp:10-73 + crt_triangle.h:25-33
bool read_meshes(const Value *list, const MaterialTable &materials, std::vector<Vertex> &vertices, std::vector<Triangle> &triangles) {
    if (!list || !list->is_array())
        return false;
    if (list->kind == Value::Kind::NumberArray)
        return list->empty();
    size_t vertex_count = 0, triangle_count = 0;
    for (const Value &obj : list->items) {
        if (!obj.is_object())
            return false;
        const Value *pos = obj.find("vertices"), *idx = obj.find("triangles");
        if (!pos || !pos->is_array() || !idx || !idx->is_array() || idx->size() % 3 != 0)
            return false;
        vertex_count += pos->size() / 3;
        triangle_count += idx->size() / 3;
    }
    vertices.reserve(vertex_count);
    triangles.reserve(triangle_count);

    for (const Value &obj : list->items) {
        const Value *pos = obj.find("vertices"), *idx = obj.find("triangles");
        const Value *mat = obj.find("material_index");
        if (!mat || !mat->is_int())
            return false;
        const int material_index = mat->as_int();
        if (material_index < 0 || size_t(material_index) >= materials.flags.size())
            return false;
        if (pos->size() % 3 != 0 || (!pos->empty() && pos->kind != Value::Kind::NumberArray))
            return false;
        if (!idx->empty() && (idx->kind != Value::Kind::NumberArray || !idx->numbers_all_int))
            return false;
        const size_t count = pos->size() / 3;
        const Value *uvs = obj.find("uvs");
        if (uvs) {
            if (!uvs->is_array() || uvs->size() % 3 != 0 || (!uvs->empty() && uvs->kind != Value::Kind::NumberArray))
                return false;
            if (uvs->size() / 3 != count)
                return false;
        }
        const size_t base = vertices.size();
        for (size_t i = 0; i < count; ++i) {
            Vertex v;
            v.position = Vector{ float(pos->numbers[3 * i]), float(pos->numbers[3 * i + 1]), float(pos->numbers[3 * i + 2]) };
            if (uvs)
                v.uv = Vector{ float(uvs->numbers[3 * i]), float(uvs->numbers[3 * i + 1]), float(uvs->numbers[3 * i + 2]) };
            vertices.push_back(v);
        }
        const TriangleFlags flags = materials.flags[size_t(material_index)];
        for (size_t i = 0; i + 2 < idx->size(); i += 3) {
            const int i0 = int(idx->numbers[i]), i1 = int(idx->numbers[i + 1]), i2 = int(idx->numbers[i + 2]);
            if (i0 < 0 || i1 < 0 || i2 < 0 || size_t(i0) >= count || size_t(i1) >= count || size_t(i2) >= count)
                return false;
            Triangle t{ int(base) + i0, int(base) + i1, int(base) + i2, {}, material_index, flags };
            const Vector e0 = vertices[size_t(t.v1)].position - vertices[size_t(t.v0)].position;
            const Vector e1 = vertices[size_t(t.v2)].position - vertices[size_t(t.v0)].position;
            t.face_normal = e0.cross(e1).normalized(); // crt_triangle.h:30-32
            // smooth normals: un-weighted sum of adjacent face normals (crt_mesh.cpp:19-25)
            vertices[size_t(t.v0)].normal += t.face_normal;
            vertices[size_t(t.v1)].normal += t.face_normal;
            vertices[size_t(t.v2)].normal += t.face_normal;
            triangles.push_back(t);
        }
        // EVERY vertex parsed so far is re-normalised after each object (crt_mesh.cpp:27-29); unreferenced
        // vertices become NaN, and normals of earlier objects drift by an ulp per pass -- reproduced on purpose.
        for (Vertex &v : vertices)
            v.normal.normalize();
    }
    return true;
}

bool read_lights(const Value *list, std::vector<Light> &out) {
    if (!list || !list->is_array())
        return false;
    if (list->kind == Value::Kind::NumberArray)
        return list->empty();
    for (const Value &v : list->items) {
        if (!v.is_object())
            return false;
        const Value *intensity = v.find("intensity");
        if (!intensity || !intensity->is_number())
            return false;
        Light l{ intensity->as_float(), {} };
        if (!read_vector(v.find("position"), l.position))
            return false;
        out.push_back(l);
    }
    return true;
}

bool read_flag(const Value &settings, const char *name, bool &value) {
    if (const Value *v = settings.find(name)) {
        if (!v->is_bool())
            return false;
        value = v->boolean;
    }
    return true;
}

} // namespace

std::optional<Scene> read_scene_from_memory(std::string_view text, const std::filesystem::path &asset_root) {
    Value doc;
    if (!jsonreader::parse(text.data(), text.size(), doc) || !doc.is_object())
        return std::nullopt;

    const Value *settings = doc.find("settings");
    if (!settings || !settings->is_object())
        return std::nullopt;
    Color background;
    if (!read_vector(settings->find("background_color"), background))
        return std::nullopt;
    const Value *camera_value = doc.find("camera");
    if (!camera_value)
        return std::nullopt;
    const Value *image_settings = settings->find("image_settings");
    if (!image_settings)
        return std::nullopt;
    std::optional<Camera> camera = read_camera(camera_value, image_settings);
    if (!camera)
        return std::nullopt;

    int bucket_size = DEFAULT_SCENE_BUCKET_SIZE;
    if (const Value *b = image_settings->find("bucket_size")) {
        if (!b->is_int())
            return std::nullopt;
        bucket_size = b->as_int();
    }

    TextureTable textures;
    if (const Value *t = doc.find("textures")) {
        TextureTable parsed;
        if (read_textures(*t, asset_root, parsed))
            textures = std::move(parsed);
    }

    MaterialTable materials;
    if (!read_materials(doc.find("materials"), textures, materials))
        return std::nullopt;

    Scene scene{ background, *camera };
    if (!read_meshes(doc.find("objects"), materials, scene.vertices, scene.triangles))
        return std::nullopt;
    // the reference builds its tree here (crt_json.cpp:606); ours is built on first use (Scene::reference_tree)

    if (!read_lights(doc.find("lights"), scene.lights))
        return std::nullopt;

    bool gi_on = false, reflections_on = true, refractions_on = true; // :616
    if (!read_flag(*settings, "gi_on", gi_on) || !read_flag(*settings, "reflections_on", reflections_on) ||
        !read_flag(*settings, "refractions_on", refractions_on))
        return std::nullopt;

    scene.textures = std::move(textures.textures);
    scene.materials = std::move(materials.materials);
    scene.bucket_size = bucket_size;
    scene.gi_on = gi_on;
    scene.reflections_on = reflections_on;
    scene.refractions_on = refractions_on;
    return scene;
}

std::optional<Scene> read_scene_from_istream(std::istream &is, const std::filesystem::path &asset_root) {
    std::string text{ std::istreambuf_iterator<char>{ is }, std::istreambuf_iterator<char>{} };
    return read_scene_from_memory(text, asset_root);
}

} // namespace crt::json
