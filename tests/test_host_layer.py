"""CPU: the C++ host layer's thread pool and work queue (Source/engine/tools) keep the reference's semantics
(thread_pool.h:14-26, work_queue.h:6-16); the Python mirror of the same two calls; every header of Source/ compiles."""
import os
import subprocess
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_thread_pool_and_work_queue_cpp(tmp_path):
    exe = str(tmp_path / "host_layer_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-pthread", "-I", os.path.join(ROOT, "Source"),
                           os.path.join(ROOT, "tests", "cpp", "host_layer_test.cpp"), "-o", exe])
    out = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stdout


def test_host_layer_headers_compile_with_the_reference_names(built, tmp_path):
    """A translation unit written the way app.cpp uses the engine (names, argument order) compiles and links against the library."""
    src = tmp_path / "uses_engine.cpp"
    src.write_text('''
#include "engine/renderer/renderer.h"
#include "engine/tools/OBJ_loader.h"
int main(int argc, char**) {
    if (argc < 100) return 0;   // compile + link only
    ThreadPool tpool; tpool.threads.allocate(8);
    Model model = {}; load_model_data(model.data, "x.obj", tpool);
    model.surrounding_aabb = get_AABB(model.data); resize_scale(model, 40.f); translate_to(model, {0.f, -15.f, -38.f});
    Scene scene; scene.gpu_devices.add(0); scene.gpu_devices.add(1);
    uint32 n; prep_scene(scene, n);
    Texture tex; Setup_Texture(tex, TextureFileType::BMP, 64, 64);
    Camera cm; RenderSettings rs{}; set_camera(cm, {0.1f, 2.f, 0.f}, {-.1f, -0.5f, -1.f}, rs, 1.0f);
    update_camera_pos(cm, {1.f, 2.f, 3.f}, {0.f, 0.f, -1.f});
    Set_Pixel(vec3b{255, 128, 0}, tex, 3, 4); transform(model, {0.f, 1.f, 0.f}, 0.5f);
    RenderInfo info; info.camera_tex = &tex; info.camera = &cm; info.scene = &scene;
    start_render_from_camera(info, tpool);
    while (wait_for_render_from_camera_to_finish(info, tpool, 33)) { info.twq.clear(); wait_for_pool(tpool, 0xffffffffu); }
    Optimized_Ray r{}; TriangleIntersectionData td{}; KD_Node* hs = nullptr; LeafNodePair* ls = nullptr;
    f32 t = get_ray_kd_tree_intersection(r, scene.models[0].kd_tree, td, &hs, ls);
    activate_pool(tpool, [](void*) {}, nullptr);
    free_scene_memory(scene); Free_Texture(tex);
    return t > tolerance;
}
''')
    exe = str(tmp_path / "uses_engine")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-pthread", "-I", os.path.join(ROOT, "Source"), "-I", os.path.join(ROOT, "include"),
                           str(src), "-o", exe, "-L", os.path.join(ROOT, "atray_b200"), "-latray_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "atray_b200")])
    assert subprocess.call([exe]) == 0


def test_python_pool_mirror(built):
    import atray_b200 as A
    assert A.wait_for_pool([], 0xFFFFFFFF) == 0          # never activated (what app.cpp:169 sees after the cancel)
    ran = []
    pool = [None] * 3
    A.activate_pool(pool, lambda d: (time.sleep(0.15), d.append(threading.get_ident())), ran)
    assert A.wait_for_pool(pool, 10) == 1                # timed out
    assert A.wait_for_pool(pool) == 0 and len(ran) == 3


def test_host_code_is_clean_under_asan_and_ubsan(meshes, tmp_path):
    """`make -C atray_b200/csrc asan`: OBJ loader (1, 3, 8 chunks), transforms, both tree builders, flattening and the camera basis
    under AddressSanitizer + UndefinedBehaviorSanitizer (-fno-sanitize-recover) on a procedural mesh, the reference's assets when
    staged, an empty file and a file of junk lines. The kernels cannot be sanitised on this pool (compute-sanitizer is closed there)."""
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "atray_b200", "csrc"), "asan"], stdout=subprocess.DEVNULL)
    junk = tmp_path / "junk.obj"
    junk.write_text("v 1 2\nv\nf 1 2 3\nvn x y z\nf 1//1 2//2\n# comment\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\nf -1 -2 -3\nf 7 8 9\n\n")
    empty = tmp_path / "empty.obj"
    empty.write_text("")
    files = [p for p, _ in meshes.values()] + [str(junk), str(empty)]
    out = subprocess.run([os.path.join(ROOT, "atray_b200", "csrc", "host_sanitize")] + files, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True, timeout=600, env=dict(os.environ, ASAN_OPTIONS="detect_leaks=1"))
    assert out.returncode == 0, out.stdout[-3000:]
    assert "ERROR: AddressSanitizer" not in out.stdout and "runtime error" not in out.stdout, out.stdout[-3000:]
