// _crt: the CPython entry point Blender's add-on calls (reference: src/python/py_crt_module.cpp).
//   _crt.render_scene_from_dict(scene: dict, asset_root: str, settings: _crt.RendererSettings)
//        -> list[(r, g, b, 1.0)]   W*H tuples, rows bottom-up                       (reference :51-119)
//   _crt.render_scene_from_dict_buffer(...) -> (width, height, bytes)   NEW fast path: float32 RGBA, rows bottom-up,
//        for callers that can take a buffer (numpy.frombuffer) instead of 8 M tuples at 4K   
//   _crt.RendererSettings((max_ray_depth, diffuse_reflection_ray_count, shadow_bias, reflection_bias,
//                          diffuse_reflection_bias, refraction_bias))                structseq, 6 fields (:16-31)
//   _crt.DEFAULT_*                                                                     (:142-157)
// Errors: ValueError("Invalid CRT Scene dict"), TypeError("Expected a RendererSettings instance"), and -- new --
// RuntimeError when no CUDA device is usable (there is no CPU path to fall back to).
#define PY_SSIZE_T_CLEAN // "y#" takes Py_ssize_t lengths
#include <Python.h>
#include <structseq.h>

#include <filesystem>
#include <stdexcept>
#include <string>
#include <vector>

#include "core/crt_json.h"
#include "core/crt_renderer.h"

namespace {

PyTypeObject *g_settings_type = nullptr;

PyStructSequence_Field g_settings_fields[] = {
    { "max_ray_depth", "Maximum recursion depth for rays" },
    { "diffuse_reflection_ray_count", "Number of rays for diffuse reflections" },
    { "shadow_bias", "Epsilon used for shadow acne avoidance" },
    { "reflection_bias", "Epsilon used for reflection acne avoidance" },
    { "diffuse_reflection_bias", "Epsilon used for diffuse reflection acne avoidance" },
    { "refraction_bias", "Epsilon used for refraction acne avoidance" },
    { nullptr, nullptr },
};
PyStructSequence_Desc g_settings_desc = { "_crt.RendererSettings", "Renderer settings used by crt_core", g_settings_fields, 6 };

struct PyRef { // owning reference
    PyObject *p;
    explicit PyRef(PyObject *o) : p(o) {}
    ~PyRef() { Py_XDECREF(p); }
    PyRef(const PyRef &) = delete;
    PyRef &operator=(const PyRef &) = delete;
    explicit operator bool() const { return p != nullptr; }
};

bool read_settings(PyObject *obj, crt::RendererSettings &out) {
    if (!PyObject_TypeCheck(obj, g_settings_type)) {
        PyErr_SetString(PyExc_TypeError, "Expected a RendererSettings instance");
        return false;
    }
    out.max_ray_depth = uint32_t(PyLong_AsUnsignedLong(PyStructSequence_GET_ITEM(obj, 0)));
    out.diffuse_reflection_ray_count = uint32_t(PyLong_AsUnsignedLong(PyStructSequence_GET_ITEM(obj, 1)));
    out.shadow_bias = float(PyFloat_AsDouble(PyStructSequence_GET_ITEM(obj, 2)));
    out.reflection_bias = float(PyFloat_AsDouble(PyStructSequence_GET_ITEM(obj, 3)));
    out.diffuse_reflection_bias = float(PyFloat_AsDouble(PyStructSequence_GET_ITEM(obj, 4)));
    out.refraction_bias = float(PyFloat_AsDouble(PyStructSequence_GET_ITEM(obj, 5)));
    return !PyErr_Occurred();
}

// shared front half: dict -> json.dumps(ensure_ascii=True) -> loader -> render
bool render_from_args(PyObject *args, std::optional<crt::Image> &image) {
    PyObject *dict_obj, *root_obj, *settings_obj;
    if (!PyArg_ParseTuple(args, "O!UO", &PyDict_Type, &dict_obj, &root_obj, &settings_obj))
        return false;
    PyRef json_module{ PyImport_ImportModule("json") };
    if (!json_module)
        return false;
    PyRef dumps{ PyObject_GetAttrString(json_module.p, "dumps") };
    if (!dumps)
        return false;
    PyRef call_args{ PyTuple_Pack(1, dict_obj) };
    PyRef kwargs{ Py_BuildValue("{sO}", "ensure_ascii", Py_True) };
    if (!call_args || !kwargs)
        return false;
    PyRef text{ PyObject_Call(dumps.p, call_args.p, kwargs.p) };
    if (!text)
        return false;
    Py_ssize_t text_size = 0, root_size = 0;
    const char *text_bytes = PyUnicode_AsUTF8AndSize(text.p, &text_size);
    const char *root_bytes = PyUnicode_AsUTF8AndSize(root_obj, &root_size);
    if (!text_bytes || !root_bytes)
        return false;
    const std::filesystem::path asset_root{ std::u8string{ root_bytes, root_bytes + root_size } };

    std::optional<crt::Scene> scene = crt::json::read_scene_from_memory(std::string_view{ text_bytes, size_t(text_size) }, asset_root);
    if (!scene) {
        PyErr_SetString(PyExc_ValueError, "Invalid CRT Scene dict");
        return false;
    }
    crt::RendererSettings settings;
    if (!read_settings(settings_obj, settings))
        return false;
    try {
        Py_BEGIN_ALLOW_THREADS // the render touches no Python state; the reference holds the GIL throughout
        try {
            image.emplace(crt::render_image(*scene, settings));
        } catch (...) {
            Py_BLOCK_THREADS
            throw;
        }
        Py_END_ALLOW_THREADS
    } catch (const std::exception &e) {
        PyErr_SetString(PyExc_RuntimeError, e.what());
        return false;
    }
    return true;
}

PyObject *render_scene_from_dict(PyObject *, PyObject *args) {
    std::optional<crt::Image> image;
    if (!render_from_args(args, image))
        return nullptr;
    const int W = image->width, H = image->height;
    PyObject *list = PyList_New(Py_ssize_t(W) * H);
    if (!list)
        return nullptr;
    PyRef one{ PyFloat_FromDouble(1.0) };
    for (int y = 0; y < H; ++y) {
        const crt::Color *row = image->buffer.data() + size_t(y) * W;
        const Py_ssize_t out_row = Py_ssize_t(H - y - 1) * W; // vertical flip (reference :102-116)
        for (int x = 0; x < W; ++x) {
            PyObject *tuple = PyTuple_New(4);
            PyObject *r = PyFloat_FromDouble(row[x].x), *g = PyFloat_FromDouble(row[x].y), *b = PyFloat_FromDouble(row[x].z);
            if (!tuple || !r || !g || !b) {
                Py_XDECREF(tuple), Py_XDECREF(r), Py_XDECREF(g), Py_XDECREF(b);
                Py_DECREF(list);
                return nullptr;
            }
            Py_INCREF(one.p);
            PyTuple_SET_ITEM(tuple, 0, r);
            PyTuple_SET_ITEM(tuple, 1, g);
            PyTuple_SET_ITEM(tuple, 2, b);
            PyTuple_SET_ITEM(tuple, 3, one.p);
            PyList_SET_ITEM(list, out_row + x, tuple);
        }
    }
    return list;
}

PyObject *render_scene_from_dict_buffer(PyObject *, PyObject *args) {
    std::optional<crt::Image> image;
    if (!render_from_args(args, image))
        return nullptr;
    const int W = image->width, H = image->height;
    // the bytes object is filled in place: one pass over the frame, no intermediate copy
    PyObject *bytes = PyBytes_FromStringAndSize(nullptr, Py_ssize_t(size_t(W) * size_t(H) * 4 * sizeof(float)));
    if (!bytes)
        return nullptr;
    float *rgba = reinterpret_cast<float *>(PyBytes_AS_STRING(bytes));
    for (int y = 0; y < H; ++y) {
        const crt::Color *row = image->buffer.data() + size_t(y) * W;
        float *dst = rgba + size_t(H - y - 1) * W * 4; // vertical flip, as the list (reference :102-116)
        for (int x = 0; x < W; ++x) {
            dst[4 * x] = row[x].x, dst[4 * x + 1] = row[x].y, dst[4 * x + 2] = row[x].z, dst[4 * x + 3] = 1.0f;
        }
    }
    return Py_BuildValue("iiN", W, H, bytes); // N: the tuple takes over the reference
}

PyMethodDef g_methods[] = {
    { "render_scene_from_dict", render_scene_from_dict, METH_VARARGS, "Render a scene dict; list of (r, g, b, 1.0), rows bottom-up" },
    { "render_scene_from_dict_buffer", render_scene_from_dict_buffer, METH_VARARGS, "Same render; (width, height, float32 RGBA bytes)" },
    { nullptr, nullptr, 0, nullptr },
};

PyModuleDef g_module = { PyModuleDef_HEAD_INIT, "_crt", nullptr, -1, g_methods, nullptr, nullptr, nullptr, nullptr };

} // namespace

PyMODINIT_FUNC PyInit__crt() {
    PyObject *m = PyModule_Create(&g_module);
    if (!m)
        return nullptr;
    bool ok = PyModule_AddIntConstant(m, "DEFAULT_SCENE_BUCKET_SIZE", crt::DEFAULT_SCENE_BUCKET_SIZE) == 0 &&
              PyModule_AddIntConstant(m, "DEFAULT_MAX_RAY_DEPTH", crt::DEFAULT_MAX_RAY_DEPTH) == 0 &&
              PyModule_AddIntConstant(m, "DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT", crt::DEFAULT_DIFFUSE_REFLECTION_RAY_COUNT) == 0;
    const std::pair<const char *, float> floats[] = {
        { "DEFAULT_SHADOW_BIAS", crt::DEFAULT_SHADOW_BIAS },
        { "DEFAULT_REFLECTION_BIAS", crt::DEFAULT_REFLECTION_BIAS },
        { "DEFAULT_DIFFUSE_REFLECTION_BIAS", crt::DEFAULT_DIFFUSE_REFLECTION_BIAS },
        { "DEFAULT_REFRACTION_BIAS", crt::DEFAULT_REFRACTION_BIAS },
    };
    for (const auto &[name, value] : floats)
        ok = ok && PyModule_AddObject(m, name, PyFloat_FromDouble(value)) == 0;
    if (ok) {
        g_settings_type = PyStructSequence_NewType(&g_settings_desc);
        ok = g_settings_type != nullptr;
    }
    if (ok) {
        Py_INCREF(g_settings_type);
        ok = PyModule_AddObject(m, "RendererSettings", reinterpret_cast<PyObject *>(g_settings_type)) == 0;
    }
    if (!ok) {
        Py_DECREF(m);
        return nullptr;
    }
    return m;
}
