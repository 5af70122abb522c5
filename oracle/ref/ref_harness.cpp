// TEST INFRASTRUCTURE ONLY (oracle/): unity translation unit that compiles the UNMODIFIED reference hot
// path from a staged copy of /root/reference/Source (one-file aabb.h layout patch, see build_ref.sh) and
// exposes it through a small C interface for tests/ and bench.py's reference arm. The product never
// links or loads this. Only tests/, __graft_entry__.smoke() and bench.py may load oracle/_ref/libatray_ref.so.
//
// What is exercised is the reference's own code: load_model_data (OBJ_loader.cpp:278-361), get_AABB /
// resize_scale / translate_to (model.h:41-152), set_camera (camera.h:36-45), prep_scene → build_KD_tree
// (renderer.cpp:264-291, kd_tree.cpp:20-288), start_render_from_camera / wait_… (renderer.cpp:403-471),
// SetRay (ray.h:24-29), get_intersection_data (renderer.cpp:34-160) and cast_ray (renderer.cpp:213-262).
#include "engine/renderer/renderer.h"
#include "engine/tools/OBJ_loader.h"
#include "pl_linux_shim.cpp"
#include "engine/renderer/renderer.cpp"
#include "engine/renderer/kd_tree.cpp"
#include "engine/renderer/scene.cpp"
#include "engine/tools/OBJ_loader.cpp"
#include "engine/tools/texture.cpp"
#include <chrono>
#include <vector>
#include <thread>

struct RefScene
{
	Scene scene;
	ThreadPool loader_pool;
	uint32 max_nodes = 0;
	// copies taken before prep_scene frees the vertex arrays (renderer.cpp:277-281)
	std::vector<vec3f> vertices, normals, tex_coords;
	std::vector<FaceVertices> faces_vertices;
	std::vector<FaceData> faces_data;
	AABB model_aabb;
	// materials of spheres/planes live here (Sphere::material / Plane::material are plain pointers, sphere.h:5-10,
	// plane.h:5-10); scene.materials must not grow after model.data.material was taken (DBuffer reallocates)
	Material extra_materials[256];
	int n_extra = 0;
};

static void make_pool(ThreadPool& tp, int n)
{
	tp.threads.allocate(n);
}

extern "C" {

// method: 0 = CENTER, 1 = SAH (kd_tree.h:13-16). rescale <= 0 → no resize_scale (as app.cpp:70 for the dragon).
void* ref_scene_create(const char* obj_path, int loader_threads, float rescale, float tx, float ty, float tz,
                       int method, unsigned cap)
{
	RefScene* rs = new RefScene();
	make_pool(rs->loader_pool, loader_threads < 1 ? 1 : loader_threads);
	Model model = {};
	load_model_data(model.data, obj_path, rs->loader_pool);
	model.surrounding_aabb = get_AABB(model.data);
	if (rescale > 0.0f) resize_scale(model, rescale);
	translate_to(model, { tx, ty, tz });
	model.kd_tree.division_method = method ? KD_Division_Method::SAH : KD_Division_Method::CENTER;
	model.kd_tree.max_no_faces_per_node = cap;

	rs->vertices.assign(model.data.vertices.front, model.data.vertices.front + model.data.vertices.size);
	rs->normals.assign(model.data.normals.front, model.data.normals.front + model.data.normals.size);
	rs->tex_coords.assign(model.data.tex_coords.front, model.data.tex_coords.front + model.data.tex_coords.size);
	rs->faces_vertices.assign(model.data.faces_vertices.front, model.data.faces_vertices.front + model.data.faces_vertices.size);
	rs->faces_data.assign(model.data.faces_data.front, model.data.faces_data.front + model.data.faces_data.size);
	rs->model_aabb = model.surrounding_aabb;

	// materials as app.cpp:91,96 (sky = material 0, model material = index 5 upstream; index 1 here)
	Material skybox = { {0.3f,0.4f,0.5f}, {0.2f,0.3f,0.4f},0.3f };
	Material model_mat = { {0.4f,0.2f,0.2f}, {0.92f,0.5f,0.0f},0.3f };
	rs->scene.materials.add(skybox);
	rs->scene.materials.add(model_mat);
	model.data.material = &rs->scene.materials[1];
	rs->scene.models.add_nocpy(model);
	prep_scene(rs->scene, rs->max_nodes);
	return rs;
}

void ref_scene_destroy(void* h)
{
	RefScene* rs = (RefScene*)h;
	free_scene_memory(rs->scene);
	rs->scene.models.clear_buffer();
	rs->loader_pool.threads.clear();
	delete rs;
}

// counts[0..7] = vertices, normals, tex_coords, faces, tree nodes, primitive refs (sum over leaves), inner nodes, max leaf size
void ref_counts(void* h, unsigned* counts)
{
	RefScene* rs = (RefScene*)h;
	KD_Tree& t = rs->scene.models[0].kd_tree;
	counts[0] = (unsigned)rs->vertices.size();
	counts[1] = (unsigned)rs->normals.size();
	counts[2] = (unsigned)rs->tex_coords.size();
	counts[3] = (unsigned)rs->faces_vertices.size();
	counts[4] = (unsigned)t.tree.length;
	unsigned refs = 0, inner = 0, maxleaf = 0;
	for (int32 i = 0; i < t.tree.length; i++)
	{
		if (t.tree[i].has_children) inner++;
		else { refs += t.tree[i].primitives.size; if (t.tree[i].primitives.size > maxleaf) maxleaf = t.tree[i].primitives.size; }
	}
	counts[5] = refs; counts[6] = inner; counts[7] = maxleaf;
}

void ref_model_dump(void* h, float* vertices, float* normals, float* tex, int* face_v, int* face_vt, int* face_vn, float* aabb6)
{
	RefScene* rs = (RefScene*)h;
	if (vertices) memcpy(vertices, rs->vertices.data(), rs->vertices.size() * 12);
	if (normals) memcpy(normals, rs->normals.data(), rs->normals.size() * 12);
	if (tex) memcpy(tex, rs->tex_coords.data(), rs->tex_coords.size() * 12);
	for (size_t i = 0; i < rs->faces_vertices.size(); i++)
		for (int j = 0; j < 3; j++)
		{
			if (face_v) face_v[3 * i + j] = rs->faces_vertices[i].vertex_indices[j];
			if (face_vt) face_vt[3 * i + j] = rs->faces_data[i].tex_coord_indices[j];
			if (face_vn) face_vn[3 * i + j] = rs->faces_data[i].vertex_normals_indices[j];
		}
	if (aabb6) memcpy(aabb6, &rs->model_aabb, 24);
}

// Flat dump of the reference-built tree: per node box (6 f32), children_start (0 = leaf), primitive count and the
// offset of its first primitive in the concatenated (node order) primitive arrays; per primitive 9 f32 + face index.
void ref_tree_dump(void* h, float* boxes, int* children_start, unsigned* prim_count, unsigned* prim_first,
                   float* tri9, unsigned* face_index)
{
	RefScene* rs = (RefScene*)h;
	KD_Tree& t = rs->scene.models[0].kd_tree;
	unsigned off = 0;
	for (int32 i = 0; i < t.tree.length; i++)
	{
		KD_Node& n = t.tree[i];
		memcpy(boxes + 6 * i, &n.aabb, 24);
		children_start[i] = n.children_start_position;
		unsigned cnt = n.has_children ? 0u : n.primitives.size;
		prim_count[i] = cnt;
		prim_first[i] = off;
		for (unsigned k = 0; k < cnt; k++)
		{
			memcpy(tri9 + 9 * (size_t)(off + k), &n.primitives.front[k].face_vertices, 36);
			face_index[off + k] = n.primitives.front[k].face_index;
		}
		off += cnt;
	}
}

static void fill_camera(Camera& cm, int W, int H, int aa, unsigned spp, int bounce, const float* eye, const float* facing, float hfov)
{
	RenderSettings rs;
	rs.anti_aliasing = aa;
	rs.resolution.x = W;
	rs.resolution.y = H;
	rs.samples_per_pixel = spp;
	rs.bounce_limit = bounce;
	set_camera(cm, { eye[0], eye[1], eye[2] }, { facing[0], facing[1], facing[2] }, rs, hfov);
}

void ref_camera(int W, int H, const float* eye, const float* facing, float hfov, float* out24)
{
	Camera cm;
	fill_camera(cm, W, H, 0, 1, 1, eye, facing, hfov);
	// aspect, camera_z, camera_x, camera_y, eye, frame_center, h_fov, half_pixel_width, half_pixel_height
	memcpy(out24, &cm.aspect_ratio, sizeof(float) * 19);
}

// Renders through the reference's own tile thread pool. Returns wall seconds of start…finished (1 ms polls).
double ref_render(void* h, int W, int H, unsigned spp, int aa, int bounce, int threads, const float* eye,
                  const float* facing, float hfov, unsigned* bgrx_out, long long* total_ray_casts)
{
	RefScene* rs = (RefScene*)h;
	Camera cm;
	fill_camera(cm, W, H, aa, spp, bounce, eye, facing, hfov);
	Texture tex;
	tex.file_type = TextureFileType::BMP;
	tex.bmb.bytes_per_pixel = 4;
	tex.bmb.width = W;
	tex.bmb.height = H;
	tex.bmb.buffer_memory = bgrx_out;
	ThreadPool tp;
	make_pool(tp, threads);
	RenderInfo info;
	info.camera_tex = &tex;
	info.camera = &cm;
	info.scene = &rs->scene;
	// renderer.cpp:385 allocates (capacity + 16) BYTES for the leaf stack: oversize it.
	info.hit_stack_capacity = rs->max_nodes + 8;
	info.leaf_stack_capacity = 16 * (rs->max_nodes + 2) + 64;
	auto t0 = std::chrono::steady_clock::now();
	start_render_from_camera(info, tp);
	while (wait_for_render_from_camera_to_finish(info, tp, 1)) {}
	auto t1 = std::chrono::steady_clock::now();
	if (total_ray_casts) *total_ray_casts = info.total_ray_casts;
	info.twq.jobs.clear();
	tp.threads.clear();
	return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"

// ---- per-ray / per-pixel dumps through the reference's own functions --------------------------------------------
struct TraceScratch
{
	RNG_Stream rng;
	DBuffer<KD_Node*> hit_stack;
	DBuffer<LeafNodePair> leaf_stack;
	RayCastTools tools;
	void* leaf_base;
	TraceScratch(uint32 max_nodes)
	{
		rng.state = 1; rng.stream = 1;
		hit_stack.capacity = max_nodes + 8;
		leaf_stack.capacity = max_nodes + 8;
		hit_stack.front = (KD_Node**)pl_buffer_alloc(hit_stack.capacity * sizeof(KD_Node*));
		leaf_base = pl_buffer_alloc((leaf_stack.capacity + 1) * sizeof(LeafNodePair));
		leaf_stack.front = (LeafNodePair*)leaf_base;
		*leaf_stack.front = { 0,-MAX_FLOAT };	// sentinel as renderer.cpp:386-387
		leaf_stack.front++;
		tools.rng_stream = &rng;
		tools.hit_stack = &hit_stack;
		tools.leaf_stack = &leaf_stack;
	}
	~TraceScratch() { pl_buffer_free(hit_stack.front); pl_buffer_free(leaf_base); }
};

static const unsigned REF_MISS = 0xffffffffu;

template <typename F>
static void parallel_rows(int n, int threads, F f)
{
	if (threads <= 1) { f(0, n, 0); return; }
	std::vector<std::thread> th;
	volatile int32 next = 0;
	for (int t = 0; t < threads; t++)
		th.emplace_back([&, t]() {
			for (;;)
			{
				int32 i = __sync_fetch_and_add(&next, 16);
				if (i >= n) break;
				f(i, i + 16 > n ? n : i + 16, t);
			}
		});
	for (auto& x : th) x.join();
}

static inline unsigned pack_bgrx(vec3f c)
{
	c = clamp(c, 0.0f, 1.0f);
	vec3b b = rgb_float_to_byte(c);
	return (unsigned)b.z | ((unsigned)b.y << 8) | ((unsigned)b.x << 16);	// Set_Pixel, texture.h:27-38
}

extern "C" {

// Primary rays with the arithmetic of renderer.cpp:317-351 (AA off), one get_intersection_data per pixel.
void ref_trace_primary(void* h, int W, int H, const float* eye, const float* facing, float hfov, int threads,
                       unsigned* face_out, float* t_out, float* uv_out, unsigned* bgrx_out)
{
	RefScene* rs = (RefScene*)h;
	Camera cm;
	fill_camera(cm, W, H, 0, 1, 1, eye, facing, hfov);
	parallel_rows(H, threads, [&](int y0, int y1, int) {
		TraceScratch sc(rs->max_nodes);
		for (int32 y = y0; y < y1; y++)
		{
			f32 film_y = -1.0f + 2.0f * ((f32)y / (f32)cm.render_settings.resolution.y);
			for (int32 x = 0; x < W; x++)
			{
				f32 film_x = (-1.0f + 2.0f * ((f32)x / (f32)cm.render_settings.resolution.x)) * cm.h_fov * cm.aspect_ratio;
				vec3f pixel_pos = cm.frame_center + (cm.camera_x * film_x) + (cm.camera_y * film_y);
				Ray ray = {};
				SetRay(ray, cm.eye, pixel_pos);
				IntersectionData id;
				get_intersection_data(ray, rs->scene, id, sc.tools);
				size_t p = (size_t)y * W + x;
				bool hit = id.type == ObjectType::TRIANGLE;
				if (face_out) face_out[p] = hit ? id.tid.face_index : REF_MISS;
				if (t_out) t_out[p] = id.distance_at_intersection;
				if (uv_out) { uv_out[2 * p] = hit ? id.tid.u : 0.f; uv_out[2 * p + 1] = hit ? id.tid.v : 0.f; }
				if (bgrx_out) bgrx_out[p] = pack_bgrx(id.hit_material->emission_color);
			}
		}
	});
}

// Arbitrary ray list: rays = n x {ox,oy,oz,dx,dy,dz} (direction used as given, as get_intersection_data does).
void ref_trace_rays(void* h, long long n, const float* rays, int threads, unsigned* face_out, float* t_out)
{
	RefScene* rs = (RefScene*)h;
	int blocks = (int)((n + 4095) / 4096);
	parallel_rows(blocks, threads, [&](int b0, int b1, int) {
		TraceScratch sc(rs->max_nodes);
		for (long long i = (long long)b0 * 4096; i < (long long)b1 * 4096 && i < n; i++)
		{
			Ray ray;
			ray.origin = { rays[6 * i], rays[6 * i + 1], rays[6 * i + 2] };
			ray.direction = { rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5] };
			IntersectionData id;
			get_intersection_data(ray, rs->scene, id, sc.tools);
			bool hit = id.type == ObjectType::TRIANGLE;
			face_out[i] = hit ? id.tid.face_index : REF_MISS;
			t_out[i] = id.distance_at_intersection;
		}
	});
}

// Same list with TriangleIntersectionData's u, v of the winning triangle (a miss leaves uv untouched).
void ref_trace_rays_uv(void* h, long long n, const float* rays, int threads, unsigned* face_out, float* t_out, float* uv_out)
{
	RefScene* rs = (RefScene*)h;
	int blocks = (int)((n + 4095) / 4096);
	parallel_rows(blocks, threads, [&](int b0, int b1, int) {
		TraceScratch sc(rs->max_nodes);
		for (long long i = (long long)b0 * 4096; i < (long long)b1 * 4096 && i < n; i++)
		{
			Ray ray;
			ray.origin = { rays[6 * i], rays[6 * i + 1], rays[6 * i + 2] };
			ray.direction = { rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5] };
			IntersectionData id;
			get_intersection_data(ray, rs->scene, id, sc.tools);
			bool hit = id.type == ObjectType::TRIANGLE;
			face_out[i] = hit ? id.tid.face_index : REF_MISS;
			t_out[i] = id.distance_at_intersection;
			if (hit) { uv_out[2 * i] = id.tid.u; uv_out[2 * i + 1] = id.tid.v; }
		}
	});
}

} // extern "C"
static inline uint64 splitmix64(uint64 x)
{
	x += 0x9e3779b97f4a7c15ULL;
	x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
	x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
	return x ^ (x >> 31);
}

extern "C" {
// Jittered (anti_aliasing = TRUE) pixel loop of renderer.cpp:334-347,358-365 with a harness-defined
// per-pixel stream (the reference seeds from rdtsc × thread id, renderer.cpp:376-378): pixel p = y*W+x uses
// RNG_Stream{ state = splitmix64(seed ^ p), stream = 2p+1 }. cast_ray is the reference's own (it draws 3
// more numbers per non-sky hit, renderer.cpp:246). Also reports per-pixel count of samples that hit geometry.
void ref_render_jitter(void* h, int W, int H, unsigned spp, int bounce, unsigned long long seed, const float* eye,
                       const float* facing, float hfov, int threads, unsigned* bgrx_out, unsigned* hit_samples_out)
{
	RefScene* rs = (RefScene*)h;
	Camera cm;
	fill_camera(cm, W, H, 1, spp, bounce, eye, facing, hfov);
	parallel_rows(H, threads, [&](int y0, int y1, int) {
		TraceScratch sc(rs->max_nodes);
		for (int32 y = y0; y < y1; y++)
		{
			f32 film_y = -1.0f + 2.0f * ((f32)y / (f32)cm.render_settings.resolution.y);
			for (int32 x = 0; x < W; x++)
			{
				uint64 p = (uint64)y * W + x;
				sc.rng.state = splitmix64(seed ^ p);
				sc.rng.stream = 2 * p + 1;
				f32 film_x = (-1.0f + 2.0f * ((f32)x / (f32)cm.render_settings.resolution.x)) * cm.h_fov * cm.aspect_ratio;
				vec3f flt_pixel_color;
				Ray ray = {};
				unsigned hits = 0;
				for (uint32 i = 0; i < spp; i++)
				{
					f32 x_off = rand_bi(sc.tools.rng_stream) * cm.half_pixel_width + film_x;
					f32 y_off = rand_bi(sc.tools.rng_stream) * cm.half_pixel_height + film_y;
					vec3f pixel_pos = cm.frame_center + (cm.camera_x * x_off) + (cm.camera_y * y_off);
					SetRay(ray, cm.eye, pixel_pos);
					int64 casts = 0;
					flt_pixel_color += cast_ray(ray, rs->scene, bounce, casts, sc.tools);
					hits += (unsigned)casts;	// bounce_limit 1: casts == 1 iff the sample hit geometry (renderer.cpp:222-231,260)
				}
				flt_pixel_color = flt_pixel_color / (f32)spp;
				bgrx_out[p] = pack_bgrx(flt_pixel_color);
				if (hit_samples_out) hit_samples_out[p] = hits;
			}
		}
	});
}

// Scene primitives as app.cpp:111-129 constructs them (the adds are commented out upstream, app.cpp:139-142).
// m7 = emission rgb, reflection rgb, scatter (material.h:4-9). Plane normals are normalised as prep_scene does
// (renderer.cpp:267-270).
int ref_scene_add_sphere(void* h, const float* center, float radius, const float* m7)
{
	RefScene* rs = (RefScene*)h;
	if (rs->n_extra >= 256) return -1;
	Material* m = &rs->extra_materials[rs->n_extra++];
	*m = { {m7[0],m7[1],m7[2]}, {m7[3],m7[4],m7[5]}, m7[6] };
	Sphere sp;
	sp.center = { center[0], center[1], center[2] };
	sp.radius = radius;
	sp.material = m;
	rs->scene.spheres.add(sp);
	return 0;
}
int ref_scene_add_plane(void* h, const float* normal, float distance, const float* m7)
{
	RefScene* rs = (RefScene*)h;
	if (rs->n_extra >= 256) return -1;
	Material* m = &rs->extra_materials[rs->n_extra++];
	*m = { {m7[0],m7[1],m7[2]}, {m7[3],m7[4],m7[5]}, m7[6] };
	Plane pl;
	pl.normal = { normal[0], normal[1], normal[2] };
	normalize(pl.normal);
	pl.distance = distance;
	pl.material = m;
	rs->scene.planes.add(pl);
	return 0;
}

// The reference's pixel loop (renderer.cpp:314-365) with its own cast_ray (213-262) at any bounce_limit, AA on or
// off, and the harness-defined per-pixel stream of ref_render_jitter (the reference seeds per thread from rdtsc).
// AA off: one SetRay, cast_ray repeated spp times on the same ray (349-356); the stream keeps running across samples.
// t_out / kind_out / prim_out describe the nearest hit of the LAST sample's primary ray (get_intersection_data on it
// once more; it draws no random numbers): kind 0 = sky, 1 = triangle, 2 = sphere, 3 = plane.
void ref_render_path(void* h, int W, int H, unsigned spp, int aa, int bounce, unsigned long long seed, const float* eye,
                     const float* facing, float hfov, int threads, unsigned* bgrx_out, float* t_out, unsigned* prim_out,
                     unsigned* kind_out, long long* casts_out)
{
	RefScene* rs = (RefScene*)h;
	Camera cm;
	fill_camera(cm, W, H, aa, spp, bounce, eye, facing, hfov);
	std::vector<long long> casts_per_thread(threads < 1 ? 1 : threads, 0);
	parallel_rows(H, threads, [&](int y0, int y1, int tid) {
		TraceScratch sc(rs->max_nodes);
		for (int32 y = y0; y < y1; y++)
		{
			f32 film_y = -1.0f + 2.0f * ((f32)y / (f32)cm.render_settings.resolution.y);
			for (int32 x = 0; x < W; x++)
			{
				uint64 p = (uint64)y * W + x;
				sc.rng.state = splitmix64(seed ^ p);
				sc.rng.stream = 2 * p + 1;
				f32 film_x = (-1.0f + 2.0f * ((f32)x / (f32)cm.render_settings.resolution.x)) * cm.h_fov * cm.aspect_ratio;
				vec3f flt_pixel_color;
				Ray ray = {};
				int64 casts = 0;
				if (aa)
				{
					for (uint32 i = 0; i < spp; i++)
					{
						f32 x_off = rand_bi(sc.tools.rng_stream) * cm.half_pixel_width + film_x;
						f32 y_off = rand_bi(sc.tools.rng_stream) * cm.half_pixel_height + film_y;
						vec3f pixel_pos = cm.frame_center + (cm.camera_x * x_off) + (cm.camera_y * y_off);
						SetRay(ray, cm.eye, pixel_pos);
						flt_pixel_color += cast_ray(ray, rs->scene, bounce, casts, sc.tools);
					}
				}
				else
				{
					vec3f pixel_pos = cm.frame_center + (cm.camera_x * film_x) + (cm.camera_y * film_y);
					SetRay(ray, cm.eye, pixel_pos);
					for (uint32 i = 0; i < spp; i++)
						flt_pixel_color += cast_ray(ray, rs->scene, bounce, casts, sc.tools);
				}
				casts_per_thread[tid] += casts;
				flt_pixel_color = flt_pixel_color / (f32)spp;
				bgrx_out[p] = pack_bgrx(flt_pixel_color);
				if (t_out || prim_out || kind_out)
				{
					IntersectionData id;
					get_intersection_data(ray, rs->scene, id, sc.tools);
					unsigned kind = id.type == ObjectType::TRIANGLE ? 1u : id.type == ObjectType::SPHERE ? 2u : id.type == ObjectType::PLANE ? 3u : 0u;
					if (t_out) t_out[p] = id.distance_at_intersection;
					if (prim_out) prim_out[p] = kind == 1u ? id.tid.face_index : REF_MISS;
					if (kind_out) kind_out[p] = kind;
				}
			}
		}
	});
	long long total = 0;
	for (long long c : casts_per_thread) total += c;
	if (casts_out) *casts_out = total;
}

// PCG32 known-answer helper: draws n values of random_u32 (PL_math.h:506-516).
void ref_pcg32(unsigned long long state, unsigned long long stream, int n, unsigned* out)
{
	RNG_Stream s; s.state = state; s.stream = stream;
	for (int i = 0; i < n; i++) out[i] = random_u32(&s);
}

// The reference's BMP writer on a caller-supplied BGRX bitmap (texture.cpp:66-115, 147-158): "<file_name>_<id>.bmp".
int ref_write_bmp(const unsigned* pixels, unsigned W, unsigned H, const char* file_name)
{
	Texture tex;
	if (!Setup_Texture(tex, TextureFileType::BMP, W, H)) return 0;
	memcpy(tex.bmb.buffer_memory, pixels, (size_t)W * H * 4);
	const bool ok = Write_To_File(tex, file_name);
	pl_buffer_free(tex.bmb.buffer_memory);
	return ok ? 1 : 0;
}

} // extern "C"
