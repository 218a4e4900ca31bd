/* pine_b200.h -- C ABI of the B200 (sm_100a) path-tracing backend for Pine's PathIntegrator.
 *
 * This is the drop-in boundary: everything `PathIntegrator::render(Scene&)`
 * (reference src/pine/impl/integrator/path.cpp:26-41, called through
 * `class Integrator { virtual void render(Scene&) }`, src/pine/core/integrator.h:20-25) does from
 * the moment the scene is complete until `Film::pixels` is filled happens behind these entry
 * points. Plain pointers and sizes only; no C++ or torch types. All functions return 0 on success
 * or a negative PB2_E_* code; `pb2_last_error()` gives the message of the calling thread's last
 * failure. Nothing here falls back to the CPU: without a CUDA device every call fails.
 *
 * Shape records carry EXACTLY the data members of the reference's shape structs (cited per
 * struct), so a Pine-side shim copies members verbatim and both sides intersect the same numbers.
 * Matrices are column-major (Pine's Matrix4<T>{x,y,z,w} column vectors, vecmath.h:556-616).
 */
#ifndef PINE_B200_H
#define PINE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB2_OK 0
#define PB2_E_INVALID -1     /* bad argument / malformed scene */
#define PB2_E_UNSUPPORTED -2 /* scene content the CUDA path does not implement (no CPU fallback) */
#define PB2_E_CUDA -3        /* CUDA runtime error or no device */
#define PB2_E_STATE -4       /* call order violated (e.g. render before upload) */

/* ---- shapes (members of the reference structs) ---------------------------------------------- */
typedef struct pb2_rect { /* pine::Rect, geometry.h:79-96 */
  float position[3], ex[3], ey[3], n[3];
  float lx, ly;
  float rx[3], ry[3];
} pb2_rect;

typedef struct pb2_aabb { /* pine::AABB, bbox.h:29-92 */
  float lower[3], upper[3];
} pb2_aabb;

typedef struct pb2_obb { /* pine::OBB, bbox.h:94-109 (PRL `Box(aabb, m)`) */
  pb2_aabb base;
  float m[16], m_inv[16]; /* column-major; m_inv is the HOST's inverse(m) (vecmath.cpp:104-131) */
} pb2_obb;

typedef struct pb2_disk { /* pine::Disk, geometry.h:44-59 */
  float position[3], n[3], u[3], v[3];
  float r;
} pb2_disk;

typedef struct pb2_cone { /* pine::Cone, geometry.h:118-139 */
  pb2_disk bottom;
  float p[3], n[3];
  float r, h, A, A2, S;
} pb2_cone;

typedef struct pb2_sphere { /* pine::Sphere, geometry.h:28-42 */
  float c[3];
  float r;
} pb2_sphere;

typedef struct pb2_triangle { /* pine::Triangle, geometry.h:98-116 */
  float v0[3], v1[3], v2[3], n[3];
} pb2_triangle;

typedef struct pb2_plane { /* pine::Plane, geometry.h:12-26; BVH::build bounds it by position +- 100 (geometry.cpp:52) */
  float position[3], n[3], u[3], v[3];
} pb2_plane;

typedef struct pb2_cylinder { /* pine::Cylinder, geometry.h:139-155: the open tube between p0 and p1. Its two Disk
                               * members both sit at p0 (geometry.h:141), so get_aabb() is the box of a sphere of
                               * radius r around p0 -- kept: that box is what the reference's BVH tests */
  float p0[3], p1[3], n[3];
  float r;
  pb2_disk bottom, top;
} pb2_cylinder;

typedef struct pb2_mesh { /* pine::Mesh, geometry.h:158-224 */
  const float* vertices; /* n_vertices x 3 */
  const uint32_t* indices; /* n_triangles x 3 */
  const float* normals;   /* n_vertices x 3 or NULL */
  const float* texcoords; /* n_vertices x 2 or NULL */
  uint32_t n_vertices, n_triangles;
} pb2_mesh;

enum pb2_shape_kind {
  PB2_SHAPE_AABB = 0,
  PB2_SHAPE_OBB = 1,
  PB2_SHAPE_SPHERE = 2,
  PB2_SHAPE_TRIANGLE = 3,
  PB2_SHAPE_CONE = 4,
  PB2_SHAPE_RECT = 5,
  PB2_SHAPE_DISK = 6,
  PB2_SHAPE_MESH = 7,
  PB2_SHAPE_PLANE = 8,
  PB2_SHAPE_CYLINDER = 9
};

/* ---- procedural material parameters: Node3f / Nodef graphs (node.h:105-250, node.cpp:8-27) ------------------
 * A graph is flattened to a POSTFIX program of pb2_node_instr, terminated by PB2_NODE_END, evaluated at the
 * shading point with NodeEvalCtx{p, n, uv} (node.h:13-19). Every value on the evaluation stack is a vec3; a
 * float-valued node (Nodef) carries its value in all three components, so the float operators are the vec3
 * operators component by component -- the same float expressions the reference evaluates. At most
 * PB2_NODE_MAX_STACK values are live. NodeNoisef / NodeNoise3f and NodeFunction (a PRL closure) have no
 * device form: a material using them is rejected (PB2_E_UNSUPPORTED), there is no CPU fallback. */
enum pb2_node_op {
  PB2_NODE_END = 0,
  PB2_NODE_CONST = 1,        /* push a[0..2]  (NodeConstant<vec3>; NodeConstant<float> v as {v, v, v}) */
  PB2_NODE_POSITION = 2,     /* push ctx.p    (node.h:116-120) */
  PB2_NODE_NORMAL = 3,       /* push ctx.n    (node.h:121-125) */
  PB2_NODE_UV = 4,           /* push vec3(ctx.uv) = {u, v, 0} (node.h:126-130) */
  PB2_NODE_ADD = 5, PB2_NODE_SUB = 6, PB2_NODE_MUL = 7, PB2_NODE_DIV = 8, /* pop b, pop a, push a op b (node.h:132-153) */
  PB2_NODE_POW = 9,          /* powf per component */
  PB2_NODE_NEG = 10, PB2_NODE_ABS = 11, PB2_NODE_SQR = 12, PB2_NODE_SQRT = 13,
  PB2_NODE_FRACT = 14,       /* x - floor(x) (node.h:155-179, psl/math.h) */
  PB2_NODE_COMPONENT = 15,   /* NodeComponent: pop x, push {x[k], x[k], x[k]}, k = int(a[0]) in 0..2 (node.h:181-193) */
  PB2_NODE_VEC3 = 16,        /* NodeToVec3(x, y, z): pop z, y, x (floats), push {x, y, z}; the one-argument form is
                              * the identity on this stack and emits nothing (node.h:195-207) */
  PB2_NODE_CHECKERBOARD = 17, /* pop p; x = fract(p) - a[0]; push float(x.x * x.y * x.z > 0) (node.cpp:16-19) */
  PB2_NODE_IMAGE = 18,       /* NodeImage: pop p, push the texel of texture int(a[0]) (1-based) at fract(p.xy)
                              * (node.cpp:20-23) */
  PB2_NODE_IMAGEF = 19       /* NodeImagef: the same texel's x, as a float (node.cpp:24-27) */
};
#define PB2_NODE_MAX_STACK 8
typedef struct pb2_node_instr {
  int32_t op; /* pb2_node_op */
  float a[3];
} pb2_node_instr;

/* ---- materials (material.h:18-131); a Node3f/Nodef parameter is a constant, an image node or a program -- */
enum pb2_material_kind {
  PB2_MAT_EMISSIVE = 0,
  PB2_MAT_DIFFUSE = 1,
  PB2_MAT_METAL = 2,
  PB2_MAT_GLOSSY = 3,
  PB2_MAT_GLASS = 4,
  PB2_MAT_UBER = 5,
  PB2_MAT_SUBSURFACE = 6
};
typedef struct pb2_material {
  int32_t kind;
  float color[3]; /* albedo, or emitted radiance for EMISSIVE */
  float roughness, ior, metallic, transmission;
  float sigma_s[3];
  /* UBER only: image nodes of a glTF material (scene_from_gltf, fileio.cpp:285-292). 1-BASED index into
   * pb2_scene_desc.textures, 0 = none (a zero-initialised material has no textures).
   * color_texture: albedo = NodeImage(NodeUV(), image) (node.cpp:20-23) replaces `color`;
   * roughness_texture: roughness = NodeComponent(NodeImage(...), 1), metallic_texture: metallic =
   * NodeComponent(NodeImage(...), 2) (node.h:181-193) replace `roughness` / `metallic`. */
  int32_t color_texture, roughness_texture, metallic_texture;
  /* Any non-emissive material: 1-BASED index of the first instruction of the parameter's program in
   * pb2_scene_desc.node_instrs, 0 = the constant above. color_node: albedo (Node3f); roughness_node; ior_node
   * (Glossy / Glass only, material.h:57-85); metallic_node / transmission_node (Uber only, material.h:87-104).
   * A program supersedes the constant and the image node of the same parameter. Emissive::color must stay a
   * constant (the light sampler would need uv at the sampled point). */
  int32_t color_node, roughness_node, ior_node, metallic_node, transmission_node;
} pb2_material;

/* pine::Image holding Array2d<vec4u8> (image.h:11; what tinygltf hands scene_from_gltf): row-major
 * texel (x, y) at rgba[4 * (x + y * width)], row 0 first. A lookup is NEAREST with wrap:
 * texel = min(int(size * fract(uv)), size - 1), value = pow(byte / 255, 2.2) per channel
 * (Image::operator[], image.cpp:7-18) -- evaluated through a 256-entry table the HOST fills with its own
 * powf, so albedo bits equal the reference's on the same machine. */
typedef struct pb2_texture {
  const uint8_t* rgba;
  int32_t width, height;
} pb2_texture;

/* One entry of Scene::geometries (scene.h:37), in the scene's order: geometry index == position. */
typedef struct pb2_geometry {
  int32_t shape_kind;  /* pb2_shape_kind */
  int32_t shape_index; /* index into the per-kind array of pb2_scene_desc */
  int32_t material;    /* index into pb2_scene_desc.materials */
} pb2_geometry;

typedef struct pb2_camera { /* pine::ThinLenCamera, camera.h:10-27 */
  float position[3];
  float c2w[9]; /* column-major mat3 (look_at(from,to) truncated, camera.cpp:10) */
  float fov2d[2];
  float len_radius, focus_distance;
  int32_t width, height; /* Film size */
  int32_t tonemapper;    /* 0 Uncharted2, 1 ACES (film.h:9-24); used by pb2_film_finalize only */
} pb2_camera;

/* One entry of UniformLightSampler::lights (lightsampler.cpp:6-10): Scene::lights in insertion order
 * (scene.cpp:19-20,29-31) followed by the environment light, if any. */
enum pb2_light_kind {
  PB2_LIGHT_AREA = 0,        /* AreaLight(geometry), light.cpp:55-69 */
  PB2_LIGHT_POINT = 1,       /* PointLight(position, color), light.cpp:11-17 */
  PB2_LIGHT_SPOT = 2,        /* SpotLight: direction normalised, cosines precomputed, light.cpp:18-45 */
  PB2_LIGHT_DIRECTIONAL = 3, /* DirectionalLight(normalize(direction), color), light.cpp:46-53 */
  PB2_LIGHT_SKY = 4          /* Sky(sun_color) as the environment light, light.cpp:71-84; must come last */
};
typedef struct pb2_light {
  int32_t kind;
  int32_t geometry;   /* AREA: index into geometries (must be Emissive); otherwise -1 */
  float position[3];  /* POINT, SPOT */
  float direction[3]; /* SPOT, DIRECTIONAL: unit vector */
  float color[3];     /* POINT, SPOT, DIRECTIONAL: radiant colour; SKY: sun_color */
  float falloff_cos, cutoff_cos; /* SPOT */
} pb2_light;

typedef struct pb2_scene_desc {
  const pb2_geometry* geometries; uint32_t n_geometries;
  const pb2_material* materials;  uint32_t n_materials;
  const pb2_aabb* aabbs;          uint32_t n_aabbs;
  const pb2_obb* obbs;            uint32_t n_obbs;
  const pb2_sphere* spheres;      uint32_t n_spheres;
  const pb2_triangle* triangles;  uint32_t n_triangles;
  const pb2_cone* cones;          uint32_t n_cones;
  const pb2_rect* rects;          uint32_t n_rects;
  const pb2_disk* disks;          uint32_t n_disks;
  const pb2_mesh* meshes;         uint32_t n_meshes;
  /* Scene::lights (scene.h:39): geometry index of every AreaLight, in scene order
   * (scene.cpp:19-20 adds one per Emissive geometry). Delta lights and Sky: see light_records. */
  const int32_t* light_geometries; uint32_t n_lights;
  pb2_camera camera;
  const pb2_texture* textures;     uint32_t n_textures; /* may be NULL / 0 */
  /* The full light list. NULL: area lights only, taken from light_geometries. Non-NULL: supersedes
   * light_geometries (whose entries must then equal the AREA records' geometries, or n_lights = 0). */
  const pb2_light* light_records;  uint32_t n_light_records;
  /* Programs of the materials' *_node parameters, back to back, each ending in PB2_NODE_END. May be NULL / 0. */
  const pb2_node_instr* node_instrs; uint32_t n_node_instrs;
  const pb2_plane* planes;         uint32_t n_planes;    /* may be NULL / 0 */
  const pb2_cylinder* cylinders;   uint32_t n_cylinders; /* may be NULL / 0 */
} pb2_scene_desc;

typedef struct pb2_render_params {
  int32_t spp;             /* BlueSobolSampler::spp() (already rounded up to a power of 2, <=256) */
  int32_t max_path_length; /* PathIntegrator ctor argument (path.cpp:9-14), > 0 */
  /* Multi-GPU sharding. Tile sharding: rows [row_begin,row_end) of the film, all samples
   * (results bit-identical to one GPU). Sample sharding: all rows, samples
   * [sample_begin,sample_end); the film then holds SUM over those samples / spp and per-pixel RNGs
   * are seeded hash(p, sample_begin) (sampler.h:286-289 takes a sample_index). Use
   * 0,height / 0,spp for a full render. */
  int32_t row_begin, row_end;
  int32_t sample_begin, sample_end;
  int32_t accel; /* 0 = wide BVH (production), 1 = reference-order binary BVH (verification) */
  int32_t flags; /* PB2_RENDER_* */
  /* Which alternative of pine::Sampler (sampler.h:203) drives get1d/get2d; 0 keeps the meaning above.
   * PB2_SAMPLER_SOBOL: SobolSampler(spp) (sampler.h:83-164, sampler.cpp:82-113), any spp in
   * [1, 65536], streams bit-identical to the reference's. PB2_SAMPLER_UNIFORM: UniformSampler(spp,
   * sampler_seed) (sampler.h:13-39) -- the reference draws from one RNG per worker thread, so its
   * stream depends on the thread schedule; here every (pixel, sample, dimension) has its own
   * counter-based draw from the same generator (rng.h:97-144): statistically equivalent only.
   * PB2_SAMPLER_HALTON: HaltonSampler(spp), digit permutations from compute_radical_inverse_permutations
   * (lowdiscrepancy.cpp:5-17) rebuilt on the host with the same generator; draws bit-identical. */
  int32_t sampler;
  int32_t sampler_seed;
  /* Interleaved row bands (fine-grained tile sharding in ONE call): with row_parts > 1 the call renders, of
   * the bands of row_band rows that [row_begin, row_end) is cut into, those whose index k has
   * k % row_parts == row_part; all other rows are left alone (zero, or kept with PB2_RENDER_KEEP_FILM).
   * row_parts <= 1 (a zero-initialised struct): every row of the range. */
  int32_t row_band, row_parts, row_part;
} pb2_render_params;
#define PB2_SAMPLER_BLUE 0
#define PB2_SAMPLER_SOBOL 1
#define PB2_SAMPLER_UNIFORM 2
#define PB2_SAMPLER_HALTON 3 /* HaltonSampler(spp) (sampler.h:40-81, sampler.cpp:20-81): bit-identical draws */
#define PB2_RENDER_SORT_BY_MATERIAL 1 /* bucket the shade queue by BxDF kind */
#define PB2_RENDER_TIME_KERNELS 2     /* fill pb2_counters.last_{trace,shadow,shade,other}_ms */
#define PB2_RENDER_KEEP_FILM 4        /* do not clear the device film first: rows outside
                                       * [row_begin,row_end) keep what earlier calls rendered
                                       * (row-band sharding issues one call per band) */

typedef struct pb2_counters {
  uint64_t closest_rays, shadow_rays, camera_samples, kernel_launches; /* cumulative */
  /* Device times of the last pb2_render*, CUDA events on its stream. The per-kernel sums are only
   * taken when PB2_RENDER_TIME_KERNELS is set (two event records around every launch). */
  double last_render_ms;
  double last_trace_ms;   /* closest-hit traversal kernels (Accel::intersect) */
  double last_shadow_ms;  /* any-hit traversal kernels (Accel::hit) */
  double last_shade_ms;   /* shading kernels */
  double last_other_ms;   /* raygen + fold/resolve + film */
  uint64_t last_closest_rays, last_shadow_rays, last_trace_launches; /* of the last render */
  double bvh_build_ms;    /* host time of the last pb2_upload_scene: BVH build + layout */
  double upload_ms;       /* host wall time of the last host->device staging of the scene */
  uint64_t scene_bytes;   /* bytes that staging copies */
  uint32_t bvh2_nodes, wide_nodes, n_primitives, build_threads;
  /* rays of the last render the wide walk could not certify and re-traced in reference order */
  uint64_t last_closest_fallbacks, last_shadow_fallbacks;
} pb2_counters;

typedef struct pb2_context pb2_context;

const char* pb2_last_error(void);
int pb2_device_count(void);
/* Replaces: RTIntegrator ctor + per-render prologue (integrator.cpp:21-35). */
int pb2_create(int device, pb2_context** out);
void pb2_destroy(pb2_context* ctx);
/* Replaces: accel.build(&scene) + light_sampler.build(&scene) (integrator.cpp:32-33,
 * bvh.cpp:453-495, lightsampler.cpp:6-10). Copies everything it needs; builds the BVH on the host. */
int pb2_upload_scene(pb2_context* ctx, const pb2_scene_desc* scene);
/* Copies the already-built scene (BVH, primitives, materials, active sampler tables) from the
 * context's pinned host staging to the device again, asynchronously on cuda_stream (NULL = the
 * context's stream). What pb2_upload_scene does after the host build; exposed so that a caller
 * (bench.py's end-to-end leg) can put the host->device transfer inside a timed region. */
int pb2_restage_scene(pb2_context* ctx, void* cuda_stream);
/* One host build for all the PROCESSES of a node (one process per GPU): the process that ran pb2_upload_scene
 * writes its built scene -- BVHs, primitive records, pools, scalar state -- to `path` (use /dev/shm: page cache,
 * no disk), and the others call pb2_import_host_scene INSTEAD of pb2_upload_scene: the file is mapped, staged to
 * the device, and from then on the context behaves as if it had built the scene (pb2_restage_scene[_part],
 * renders); only the tree dumps pb2_bvh_info / pb2_bvh_nodes stay with the builder. The file may be unlinked
 * once every importer has returned. (Inside one process pb2_multi shares the build without a file.) */
int pb2_export_host_scene(pb2_context* ctx, const char* path);
int pb2_import_host_scene(pb2_context* ctx, const char* path);
/* Staging for REPLICAS on one NVLink domain (one context per GPU, same scene on each): instead of every GPU
 * pulling the whole scene through PCIe, part `part` of `n_parts` copies only its 1/n_parts slice of every large
 * array (>= 1 MiB) from pinned host memory -- small arrays and the sampler tables whole -- and the caller
 * all-gathers each large array IN PLACE over NVLink (NCCL: sendbuf = device_ptr + part * part_bytes, recvbuf =
 * device_ptr, count = part_bytes; pine_b200/multigpu.py restage_sharded). `arrays` receives the large arrays
 * (each padded to n_parts * part_bytes); at most max_arrays, the count in *n_arrays. */
typedef struct pb2_device_array {
  void* device_ptr;
  uint64_t bytes;      /* n_parts * part_bytes (allocation, >= the array's own size) */
  uint64_t part_bytes; /* multiple of 256 */
} pb2_device_array;
int pb2_restage_scene_part(pb2_context* ctx, void* cuda_stream, int part, int n_parts,
                           pb2_device_array* arrays, int max_arrays, int* n_arrays);
/* The Heitz et al. tables Pine links in (src/contrib/bluesobol/bluenoise_<spp>spp.cpp:2-12), as u8:
 * sobol[256*256], scramble[128*128*8], rank[128*128*8]. */
int pb2_upload_sampler_tables(pb2_context* ctx, int spp_pow2, const uint8_t* sobol,
                              const uint8_t* scramble, const uint8_t* rank);
/* Same, read from the blob pine_b200/tools/extract_tables.cpp extracts from those sources (path NULL or "" =
 * pine_b200/data/bluesobol_tables.bin beside the library); what a C++ caller without the arrays uses. */
int pb2_load_sampler_tables_file(pb2_context* ctx, const char* path, int spp_pow2);
/* Replaces: PathIntegrator::render's pixel loop (path.cpp:31-40). film_rgba: caller-owned HOST
 * buffer of width*height*4 floats, row-major p.x + p.y*width (array.h:51-55), pre-tonemap. */
int pb2_render(pb2_context* ctx, const pb2_render_params* params, float* film_rgba_host);
/* Replaces: `if (p.x % 64 == 0) set_progress(progress_2d(p, film.size()))` in the pixel loop (path.cpp:39;
 * Atomic<float> g_progress, integrator.cpp:11-19, polled by the CLI's progress thread). While a render runs, `fn`
 * is called with the fraction of its pixel waves that have COMPLETED on the device (monotone, ends at 1), from a
 * CUDA driver thread (cudaLaunchHostFunc behind each wave): it must be thread-safe, return quickly and make no CUDA
 * call -- storing an atomic, which is what pine::set_progress does, is the intended use. fn NULL clears. */
typedef void (*pb2_progress_fn)(float fraction, void* user);
int pb2_set_progress_callback(pb2_context* ctx, pb2_progress_fn fn, void* user);
/* Same, but leaves the film in DEVICE memory (for the multi-GPU NCCL reduce); pointer stays owned
 * by ctx and valid until the next render/destroy. cuda_stream: a cudaStream_t or NULL. */
int pb2_render_device(pb2_context* ctx, const pb2_render_params* params, void* cuda_stream,
                      float** film_rgba_device);
/* Film::finalize + to_uint8 (film.cpp:21-27, color.cpp:6-23, fileio.cpp:42-54) on the device film:
 * tone-map, gamma 1/2.2, *256 clamp, optional y-flip. rgba8_host: width*height*4 bytes. */
int pb2_film_finalize(pb2_context* ctx, const float* film_rgba_host, int flip_y,
                      uint8_t* rgba8_host, float* tonemapped_rgba_host);
/* The same epilogue on the film the last pb2_render* LEFT ON THE DEVICE: no upload, and 4 bytes per pixel come
 * back instead of 16 (what Film::save needs: save_film_as_image, film.cpp:12-15). tonemapper: 0 Uncharted2,
 * 1 ACES, -1 the uploaded camera's. Either output may be NULL. */
int pb2_film_finalize_device(pb2_context* ctx, int tonemapper, int flip_y, uint8_t* rgba8_host,
                             float* tonemapped_rgba_host);
/* Test hooks == Accel::intersect / Accel::hit (accel.h:16-24) on a caller ray set.
 * rays: n x {o.xyz, d.xyz, tmin, tmax}. geom/prim = -1 on miss, t = ray.tmax after traversal.
 * surf (optional, n x 8) = p, n, uv from compute_surface_info. accel as in pb2_render_params. */
int pb2_trace_closest(pb2_context* ctx, const float* rays_host, size_t n, int accel,
                      int32_t* geom, int32_t* prim, float* t, float* surf);
int pb2_trace_any(pb2_context* ctx, const float* rays_host, size_t n, int accel,
                  uint8_t* occluded);
/* Device-resident variant for benchmarking the traversal kernel alone: rays stay in HBM. */
int pb2_trace_closest_device(pb2_context* ctx, const float* rays_dev, size_t n, int accel,
                             int32_t* geom_dev, int32_t* prim_dev, float* t_dev,
                             void* cuda_stream);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * pb2_trace_closest_device with its scratch (the persistent walk's claim counter, fallback count and fallback
 * list, which the entry point above reserves with cudaMalloc as n grows) in a file-scope static __device__ array
 * of PB2_NOALLOC_MAX_RAYS + 64 ints = 4 MiB + 256 B, which every GPU that loads the library holds until exit.
 * Same arguments, same results bit for bit; the caller passes device inputs and outputs. Calls no cudaMalloc,
 * cudaFree or allocating library routine. n > PB2_NOALLOC_MAX_RAYS => PB2_E_INVALID (split the batch). The
 * first call on a device looks the array's address up (cudaGetSymbolAddress; n = 0 does only that, so make that
 * call once outside any capture). Every call on a device shares the array and it keeps the last call's contents:
 * the two counters are cleared on the launch stream before each launch, but two calls that OVERLAP on the GPU
 * (different streams with no ordering between them) race on it -- order them, or use the entry point above. */
#define PB2_NOALLOC_MAX_RAYS (1 << 20)
int pb2_trace_closest_device_noalloc(pb2_context* ctx, const float* rays_dev, size_t n, int accel,
                                     int32_t* geom_dev, int32_t* prim_dev, float* t_dev,
                                     void* cuda_stream);
/* Known-answer hooks for the sampler/camera kernels (Sampler wrapper sampler.h:286-309,
 * BlueSobolSampler sampler.h:166-201, ThinLenCamera::gen_ray camera.cpp:22-33). */
/* Test hook: every ray a render traces. Set before pb2_render* (NULL clears). The render must fit
 * one wave (small films); it appends, per depth, the closest-hit queue with its results and the
 * shadow queue with its occlusion results, up to `capacity` rays per kind (more => PB2_E_INVALID).
 * Rays are n x {o.xyz, d.xyz, tmin, tmax}, exactly what Accel::intersect / Accel::hit received. */
typedef struct pb2_ray_record {
  uint64_t capacity;
  uint64_t n_closest, n_shadow; /* out */
  float* closest_rays; float* closest_t; int32_t* closest_geom; int32_t* closest_prim;
  float* shadow_rays; uint8_t* shadow_occluded;
} pb2_ray_record;
int pb2_set_ray_record(pb2_context* ctx, pb2_ray_record* rec);
int pb2_test_sampler(pb2_context* ctx, int spp, int px, int py, int sample_index, int n_rand,
                     float* out_rand, const char* pattern, int n_samples, float* out_draws);
/* get1d/get2d draws of any sampler kind: Sampler::init(film size of the uploaded scene),
 * start_pixel((px, py), sample_index), then per sample the pattern of '1'/'2' draws. */
int pb2_test_sampler_kind(pb2_context* ctx, int sampler, int spp, int sampler_seed, int px, int py,
                          int sample_index, const char* pattern, int n_samples, float* out_draws);
int pb2_test_primary_rays(pb2_context* ctx, int sample_index, float* rays_host);
/* Known-answer hook for the node programs: Mnode<T>::operator()(NodeEvalCtx{p, n, uv}) (node.h:13-19, 265-295) of
 * the uploaded scene's program starting at `start` (1-based, as in pb2_material.*_node) at n points.
 * ctx8: n x {p.xyz, n.xyz, uv}; out3: n x vec3 (a float node's value in all three components). */
int pb2_test_eval_node(pb2_context* ctx, int start, size_t n, const float* ctx8_host, float* out3_host);
int pb2_get_counters(pb2_context* ctx, pb2_counters* out);
/* Measurement hook (SURVEY 8(d) asks for a measured L2 peak beside the HBM peak): sustained read bandwidth in
 * GB/s of a `bytes`-sized device buffer read `iters` times by a full grid with L1 bypassed (ld.global.cg).
 * bytes << 126 MB: the L2; bytes >> 126 MB: HBM. */
int pb2_measure_read_gbs(pb2_context* ctx, size_t bytes, int iters, double* gbs);
/* Debug hook (meaningful in -DPB2_COUNT tuning builds only; zeros otherwise): wide-walk node and
 * primitive tests since the last call, see pine_b200/csrc/api.cu. */
int pb2_debug_walk_stats(pb2_context* ctx, uint64_t* out10);
/* Dump of the host-built reference-order binary tree (which=-1 top level, k = k-th mesh tree) in
 * the layout of oracle/ref_harness.cpp:ref_bvh_nodes, to pin the builder node for node. */
int pb2_bvh_info(pb2_context* ctx, int which, int32_t* n_nodes, int32_t* root,
                 int32_t* n_leaf_prims);
int pb2_bvh_nodes(pb2_context* ctx, int which, float* aabbs, int32_t* children,
                  int32_t* leaf_range, int32_t* leaf_prims);

/* ---- every GPU of the box behind ONE integrator (BASELINE.json north_star; the reference interface stays
 * `Integrator::render(Scene&)`, integrator.h:20-25, registered at program_context.cpp:76-81) -----------------
 * A pb2_multi owns one pb2_context per device. One host thread calls in; the BVH is built once on the host,
 * the scene is replicated (each GPU pulls 1/N of every large array through PCIe, the rest arrives by in-place
 * ncclAllGather over NVLink), the (pixel, sample) work is sharded, and ONE ncclReduce(sum) of the float film
 * brings the result to the first device. NCCL is loaded at run time (libnccl.so.2) and only when N > 1.
 *   PB2_SHARD_SAMPLES: GPU i renders a contiguous share of [sample_begin, sample_end) of every pixel; per-pixel
 *     RNG streams are re-seeded hash(p, first sample of the share) (sampler.h:286-289 takes a sample_index), so
 *     the film equals the single-GPU film statistically, not bit for bit.
 *   PB2_SHARD_ROWS: bands of PB2_SHARD_ROW_BAND film rows dealt round-robin; every pixel has one writer and
 *     the other GPUs add zeros: BIT-IDENTICAL to the single-GPU film. */
typedef struct pb2_multi pb2_multi;
#define PB2_SHARD_SAMPLES 0
#define PB2_SHARD_ROWS 1
#define PB2_SHARD_ROW_BAND 16
/* devices NULL or n_devices <= 0: every visible CUDA device. */
int pb2_create_multi(const int* devices, int n_devices, pb2_multi** out);
void pb2_destroy_multi(pb2_multi* m);
int pb2_multi_size(pb2_multi* m);
/* The i-th device's context (owned by m), for the single-device entry points above (test hooks, counters). */
pb2_context* pb2_multi_context(pb2_multi* m, int i);
int pb2_multi_upload_sampler_tables(pb2_multi* m, int spp_pow2, const uint8_t* sobol, const uint8_t* scramble,
                                    const uint8_t* rank);
int pb2_multi_load_sampler_tables_file(pb2_multi* m, const char* path, int spp_pow2);
/* pb2_upload_scene for all devices: one host build, sharded staging + all-gather. */
int pb2_multi_upload_scene(pb2_multi* m, const pb2_scene_desc* scene);
/* Repeats only the staging (host -> devices), synchronously: the end-to-end timing leg. */
int pb2_multi_restage_scene(pb2_multi* m);
/* pb2_render over all devices; film_rgba_host as in pb2_render. Blocking. */
int pb2_multi_render(pb2_multi* m, const pb2_render_params* params, int shard_mode, float* film_rgba_host);
/* pb2_set_progress_callback for the whole job: `fn` receives the mean of the devices' fractions (called from the
 * devices' driver threads, possibly concurrently). */
int pb2_multi_set_progress_callback(pb2_multi* m, pb2_progress_fn fn, void* user);
/* pb2_film_finalize_device on the reduced film (first device) of the last pb2_multi_render. */
int pb2_multi_film_finalize(pb2_multi* m, int tonemapper, int flip_y, uint8_t* rgba8_host,
                            float* tonemapped_rgba_host);
/* Counters summed over the devices (last_render_ms: the slowest device's); optional host wall times of the
 * last film reduce and the last staging, in ms. */
int pb2_multi_get_counters(pb2_multi* m, pb2_counters* out, double* reduce_ms, double* stage_ms);

#ifdef __cplusplus
}
#endif
#endif /* PINE_B200_H */
