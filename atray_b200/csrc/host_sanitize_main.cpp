// Driver of the `asan` make target: the host half of the C ABI (OBJ loader, transforms, tree build, flattening, camera basis —
// host_model.cpp, what atray_model_* / atray_scene_upload run before anything touches the GPU) under AddressSanitizer and
// UndefinedBehaviorSanitizer. The device half cannot be sanitised on this pool (compute-sanitizer is closed on the GPU boxes).
//   host_sanitize model.obj [more.obj ...]   → exit 0 and one line per file, or the sanitizer's report
#include <cstdio>
#include <string>
#include "host_model.h"

using namespace atray;

int main(int argc, char** argv)
{
    for (int i = 1; i < argc; ++i)
        for (int threads : {1, 3, 8})
            for (int method : {0, 1})
            {
                HostModel m;
                std::string err;
                if (load_obj(argv[i], threads, m, err) != 0) { std::printf("%s: %s\n", argv[i], err.c_str()); continue; }
                compute_aabb(m);
                resize_scale(m, 40.0f);
                translate_to(m, Vec3{0.0f, -15.0f, -38.0f});
                const uint32_t cap = method ? 300u : 32u;
                if (build_tree(m, method, cap, err, threads) != 0) { std::printf("%s: build_tree: %s\n", argv[i], err.c_str()); continue; }
                FlatScene f;
                flatten(m, f);
                const CameraBasis cb = make_camera(1280, 720, Vec3{0.1f, 2.0f, 0.0f}, Vec3{-0.1f, -0.5f, -1.0f}, 1.0f);
                std::printf("%s threads %d method %d: %zu faces, %u nodes, %u slots, %u leaves, aspect %.3f\n", argv[i], threads, method,
                            m.faces.size(), f.n_nodes, f.n_slots, f.n_leaves, (double)cb.aspect);
            }
    return 0;
}
