// ACMMP_dropin.cpp - the drop-in proof: the reference's class ACMMP, RunJBU and the free helpers of ACMMP.cpp /
// ACMMP.cu re-implemented over the C ABI of include/acmmp_b200.h, against the reference's UNCHANGED header
// (/root/reference/src/ACMMP.h, found on the include path - never copied).  Linked with the reference's unchanged
// acmmp_definitions.cpp (ProcessProblem, JointBilateralUpsampling, RunFusion) and a restatement of its main()
// (oracle/dropin/dropin_main.cpp) this is the reference PROGRAM with libacmmp_b200.so underneath:
//
//     ACMMP.h, acmmp_definitions.h, acmmp_definitions.cpp    the reference's, as they lie
//     ACMMP.cpp, ACMMP.cu                                     replaced by THIS file (the INTEGRATION.md step)
//
// The header is not touched, so the class keeps its private members; the one thing the product needs per object
// (the acmmp_ctx handle) lives in a side table keyed by `this`.  TEST INFRASTRUCTURE (tests/test_gpu_dropin.py):
// built by `make dropin` into oracle/_ref/, only where /root/reference exists.  OpenCV = oracle/ref_stub.
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <sys/stat.h>

#include <mutex>
#include <unordered_map>

#include "ACMMP.h"
#include "acmmp_b200.h"

static_assert(sizeof(Camera) == sizeof(acmmp_camera), "Camera is the C ABI's acmmp_camera");

// what the reference seeds with clock64() (ACMMP.cu:624)
static unsigned long long g_seed = 0, g_runs = 0;
extern "C" void acmmp_dropin_set_seed(unsigned long long seed) {
    g_seed = seed;
    g_runs = 0;
}

namespace {

std::unordered_map<const ACMMP *, acmmp_ctx *> g_ctx;
std::mutex g_mutex;

acmmp_ctx *&ctx_of(const ACMMP *self) {
    std::lock_guard<std::mutex> lock(g_mutex);
    return g_ctx[self];
}

[[noreturn]] void die(const char *what, acmmp_ctx *ctx) {   // CudaSafeCall prints and exits (ACMMP.cpp:67-100)
    fprintf(stderr, "acmmp_b200: %s: %s\n", what, acmmp_last_error(ctx));
    exit(EXIT_FAILURE);
}
#define CK(ctx, call)                     \
    do {                                  \
        if ((call) != 0) die(#call, ctx); \
    } while (0)

std::string numbered(const std::string &folder, int id, const char *suffix) {
    char buf[64];
    snprintf(buf, sizeof buf, "/%08d%s", id, suffix);
    return folder + buf;
}

// <output_folder>/2333_%08d: one folder per reference image (acmmp_definitions.cpp:392-401)
std::string result_dir(const std::string &output_folder, int id) {
    char buf[64];
    snprintf(buf, sizeof buf, "/2333_%08d", id);
    return output_folder + buf;
}

// .dmb: int32 type (1), h, w, channels, float payload (ACMMP.cpp:263-380)
int read_dmb(const std::string &path, cv::Mat &out, int channels) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) {
        std::cout << "Error opening file " << path << std::endl;
        return -1;
    }
    int32_t hdr[4] = {0, 0, 0, 0};
    int rc = -1;
    if (fread(hdr, sizeof(int32_t), 4, f) == 4 && hdr[0] == 1 && hdr[3] == channels && hdr[1] > 0 && hdr[2] > 0) {
        out = cv::Mat(hdr[1], hdr[2], channels == 3 ? CV_32FC3 : CV_32FC1);
        const size_t n = (size_t)hdr[1] * hdr[2] * channels;
        rc = fread(out.data, sizeof(float), n, f) == n ? 0 : -1;
    }
    fclose(f);
    return rc;
}
int write_dmb(const std::string &path, const cv::Mat &m, int channels) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) {
        std::cout << "Error opening file " << path << std::endl;
        return -1;
    }
    const int32_t hdr[4] = {1, m.rows, m.cols, channels};
    fwrite(hdr, sizeof(int32_t), 4, f);
    fwrite(m.data, sizeof(float), (size_t)m.rows * m.cols * channels, f);
    fclose(f);
    return 0;
}

}  // namespace

// ---- free helpers of ACMMP.cpp that the unchanged callers use -----------------------------------------------
void CudaSafeCall(const cudaError_t error, const std::string &file, const int line) {
    if (error != cudaSuccess) {
        fprintf(stderr, "cuda error %s:%d: %s\n", file.c_str(), line, cudaGetErrorString(error));
        exit(EXIT_FAILURE);
    }
}
void CudaCheckError(const char *file, const int line) { CudaSafeCall(cudaGetLastError(), file, line); }

int readDepthDmb(const std::string file_path, cv::Mat_<float> &depth) {
    cv::Mat m;
    const int rc = read_dmb(file_path, m, 1);
    if (rc == 0) depth = m;
    return rc;
}
int readNormalDmb(const std::string file_path, cv::Mat_<cv::Vec3f> &normal) {
    cv::Mat m;
    const int rc = read_dmb(file_path, m, 3);
    if (rc == 0) normal = m;
    return rc;
}
int writeDepthDmb(const std::string file_path, const cv::Mat_<float> depth) { return write_dmb(file_path, depth, 1); }
int writeNormalDmb(const std::string file_path, const cv::Mat_<cv::Vec3f> normal) { return write_dmb(file_path, normal, 3); }

Camera ReadCamera(const std::string &cam_path) {   // cams/%08d_cam.txt (ACMMP.cpp:149-177)
    Camera camera;
    std::ifstream file(cam_path);
    std::string word;
    file >> word;   // "extrinsic"
    for (int r = 0; r < 3; ++r) file >> camera.R[3 * r] >> camera.R[3 * r + 1] >> camera.R[3 * r + 2] >> camera.t[r];
    float last_row[4];
    file >> last_row[0] >> last_row[1] >> last_row[2] >> last_row[3];
    file >> word;   // "intrinsic"
    for (int r = 0; r < 3; ++r) file >> camera.K[3 * r] >> camera.K[3 * r + 1] >> camera.K[3 * r + 2];
    float interval, planes;
    file >> camera.depth_min >> interval >> planes >> camera.depth_max;
    return camera;
}

// colour image at the size of the depth map, intrinsics rescaled with it (ACMMP.cpp:179-203)
void RescaleImageAndCamera(cv::Mat_<cv::Vec3b> &src, cv::Mat_<cv::Vec3b> &dst, cv::Mat_<float> &depth, Camera &camera) {
    const int cols = depth.cols, rows = depth.rows;
    if (cols == src.cols && rows == src.rows) {
        dst = src.clone();
        return;
    }
    const float sx = cols / static_cast<float>(src.cols), sy = rows / static_cast<float>(src.rows);
    cv::resize(src, dst, cv::Size(cols, rows), 0, 0, cv::INTER_LINEAR);
    camera.K[0] *= sx;
    camera.K[2] *= sx;
    camera.K[4] *= sy;
    camera.K[5] *= sy;
    camera.width = cols;
    camera.height = rows;
}

// pixel + depth -> world point: X_c = depth K^-1 (x, y, 1), X_w = R^T X_c + C (ACMMP.cpp:205-231)
float3 Get3DPointonWorld(const int x, const int y, const float depth, const Camera camera) {
    const float cx = depth * (x - camera.K[2]) / camera.K[0];
    const float cy = depth * (y - camera.K[5]) / camera.K[4];
    const float cz = depth;
    float3 rot;
    rot.x = camera.R[0] * cx + camera.R[3] * cy + camera.R[6] * cz;
    rot.y = camera.R[1] * cx + camera.R[4] * cy + camera.R[7] * cz;
    rot.z = camera.R[2] * cx + camera.R[5] * cy + camera.R[8] * cz;
    float3 centre;
    centre.x = -(camera.R[0] * camera.t[0] + camera.R[3] * camera.t[1] + camera.R[6] * camera.t[2]);
    centre.y = -(camera.R[1] * camera.t[0] + camera.R[4] * camera.t[1] + camera.R[7] * camera.t[2]);
    centre.z = -(camera.R[2] * camera.t[0] + camera.R[5] * camera.t[1] + camera.R[8] * camera.t[2]);
    return make_float3(rot.x + centre.x, rot.y + centre.y, rot.z + centre.z);
}

// world point -> pixel + depth in `camera` (ACMMP.cpp:244-254)
void ProjectonCamera(const float3 PointX, const Camera camera, float2 &point, float &depth) {
    float3 c;
    c.x = camera.R[0] * PointX.x + camera.R[1] * PointX.y + camera.R[2] * PointX.z + camera.t[0];
    c.y = camera.R[3] * PointX.x + camera.R[4] * PointX.y + camera.R[5] * PointX.z + camera.t[1];
    c.z = camera.R[6] * PointX.x + camera.R[7] * PointX.y + camera.R[8] * PointX.z + camera.t[2];
    depth = camera.K[6] * c.x + camera.K[7] * c.y + camera.K[8] * c.z;
    point.x = (camera.K[0] * c.x + camera.K[1] * c.y + camera.K[2] * c.z) / depth;
    point.y = (camera.K[3] * c.x + camera.K[4] * c.y + camera.K[5] * c.z) / depth;
}

float GetAngle(const cv::Vec3f &v1, const cv::Vec3f &v2) {   // ACMMP.cpp:256-261: NaN (|dot| > 1) counts as 0
    const float angle = acosf(v1[0] * v2[0] + v1[1] * v2[1] + v1[2] * v2[2]);
    return angle != angle ? 0.0f : angle;
}

// binary little-endian PLY: x y z nx ny nz (float) red green blue (uchar); colour is stored b, g, r (ACMMP.cpp:382-435)
void StoreColorPlyFileBinaryPointCloud(const std::string &plyFilePath, const std::vector<PointList> &pc) {
    std::cout << "store 3D points to ply file" << std::endl;
    FILE *f = fopen(plyFilePath.c_str(), "wb");
    if (!f) return;
    fprintf(f, "ply\nformat binary_little_endian 1.0\nelement vertex %zu\n", pc.size());
    fprintf(f, "property float x\nproperty float y\nproperty float z\nproperty float nx\nproperty float ny\nproperty float nz\n");
    fprintf(f, "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n");
    for (const PointList &p : pc) {
        float rec[6] = {p.coord.x, p.coord.y, p.coord.z, p.normal.x, p.normal.y, p.normal.z};
        if (!(rec[0] < FLT_MAX && rec[0] > -FLT_MAX) || !(rec[1] < FLT_MAX && rec[1] > -FLT_MAX) || !(rec[2] < FLT_MAX && rec[2] >= -FLT_MAX))
            rec[0] = rec[1] = rec[2] = 0.0f;
        const char rgb[3] = {(char)(int)p.color.z, (char)(int)p.color.y, (char)(int)p.color.x};
        fwrite(rec, sizeof(float), 6, f);
        fwrite(rgb, 1, 3, f);
    }
    fclose(f);
}

// RunJBU (ACMMP.cpp:1022-1087) = acmmp_jbu_depth + the depths.dmb it writes; JBU_cu runs inside the library
void RunJBU(const cv::Mat_<float> &scaled_image_float, const cv::Mat_<float> &src_depthmap, const std::string &dense_folder,
            const std::string &output_folder, const Problem &problem) {
    (void)dense_folder;
    cv::Mat_<float> depthmap(scaled_image_float.rows, scaled_image_float.cols);
    const int rc = acmmp_jbu_depth(0, scaled_image_float.ptr<float>(), scaled_image_float.cols, scaled_image_float.rows,
                                   src_depthmap.ptr<float>(), src_depthmap.cols, src_depthmap.rows, depthmap.ptr<float>());
    if (rc < 0) die("acmmp_jbu_depth", nullptr);
    if (rc > 0) {   // integer scale 1: the reference returns without writing
        std::cout << "Image.rows = Depthmap.rows" << std::endl;
        return;
    }
    const std::string result_folder = result_dir(output_folder, problem.ref_image_id);
    mkdir(result_folder.c_str(), 0777);
    writeDepthDmb(result_folder + "/depths.dmb", depthmap);
}

// class JBU is declared by the unchanged header; its device half lives in the library (jbu_depth_kernel)
JBU::JBU() : depth_h(nullptr), depth_d(nullptr), jt_d(nullptr), jp_d(nullptr) {}
JBU::~JBU() {}
void JBU::InitializeParameters(int) {}
void JBU::CudaRun() {}

// ---- class ACMMP (ACMMP.h:58-124, header unchanged) --------------------------------------------------------
ACMMP::ACMMP()
    : num_images(0), plane_hypotheses_host(nullptr), scaled_plane_hypotheses_host(nullptr), costs_host(nullptr), pre_costs_host(nullptr),
      prior_planes_host(nullptr), plane_masks_host(nullptr), sample_prior_cuda(nullptr), cameras_cuda(nullptr), texture_objects_cuda(nullptr),
      texture_depths_cuda(nullptr), plane_hypotheses_cuda(nullptr), scaled_plane_hypotheses_cuda(nullptr), costs_cuda(nullptr),
      pre_costs_cuda(nullptr), rand_states_cuda(nullptr), selected_views_cuda(nullptr), depths_cuda(nullptr), prior_planes_cuda(nullptr),
      plane_masks_cuda(nullptr) {
    acmmp_ctx *ctx = nullptr;
    if (acmmp_create(&ctx, 0) != 0) die("acmmp_create", nullptr);
    ctx_of(this) = ctx;
}

ACMMP::~ACMMP() {
    delete[] plane_hypotheses_host;
    delete[] costs_host;
    delete[] prior_planes_host;
    delete[] plane_masks_host;
    std::lock_guard<std::mutex> lock(g_mutex);
    auto it = g_ctx.find(this);
    if (it != g_ctx.end()) {
        acmmp_destroy(it->second);
        g_ctx.erase(it);
    }
}

void ACMMP::SetGeomConsistencyParams(bool multi_geometry = false) {   // ACMMP.cpp:447-454
    params.geom_consistency = true;
    params.max_iterations = 2;
    if (multi_geometry) params.multi_geometry = true;
    CK(ctx_of(this), acmmp_set_geom_consistency(ctx_of(this), multi_geometry ? 1 : 0));
}
void ACMMP::SetHierarchyParams() {
    params.hierarchy = true;
    CK(ctx_of(this), acmmp_set_hierarchy(ctx_of(this)));
}
void ACMMP::SetPlanarPriorParams() {
    params.planar_prior = true;
    CK(ctx_of(this), acmmp_set_planar_prior_params(ctx_of(this)));
}

// fixed per-pixel hypotheses of the seeded mode (ACMMP.cpp:476-484); called after CudaSpaceInitialization
// (acmmp_definitions.cpp:270-281)
void ACMMP::SetPlanarPrior(std::unique_ptr<float4> prior) {
    CK(ctx_of(this), acmmp_set_seeded_planes(ctx_of(this), reinterpret_cast<const float *>(prior.get())));
    params.seeded = true;
}

// one view as InputInitialization reads it (ACMMP.cpp:534-598): grey 8-bit -> float, camera, resize to the scale's size
static void load_view(const std::string &dense_folder, int image_id, int max_image_size, cv::Mat &image, Camera &camera) {
    cv::Mat_<uint8_t> grey = cv::imread(numbered(dense_folder + "/images", image_id, ".jpg"), cv::IMREAD_GRAYSCALE);
    grey.convertTo(image, CV_32FC1);
    camera = ReadCamera(numbered(dense_folder + "/cams", image_id, "_cam.txt"));
    camera.height = image.rows;
    camera.width = image.cols;
    if (image.cols <= max_image_size && image.rows <= max_image_size) return;
    const float fx = static_cast<float>(max_image_size) / image.cols, fy = static_cast<float>(max_image_size) / image.rows;
    const float f = std::min(fx, fy);
    const int new_cols = std::round(image.cols * f), new_rows = std::round(image.rows * f);
    const float sx = new_cols / static_cast<float>(image.cols), sy = new_rows / static_cast<float>(image.rows);
    cv::Mat_<float> scaled;
    cv::resize(image, scaled, cv::Size(new_cols, new_rows), 0, 0, cv::INTER_LINEAR);
    image = scaled.clone();
    camera.K[0] *= sx;
    camera.K[2] *= sx;
    camera.K[4] *= sy;
    camera.K[5] *= sy;
    camera.height = new_rows;
    camera.width = new_cols;
}

void ACMMP::InputInitialization(const std::string &dense_folder, const std::string &output_folder, const std::vector<Problem> &problem,
                                const int idx) {
    images.clear();
    cameras.clear();
    const Problem &local = problem[idx];
    std::vector<int> ids(1, local.ref_image_id);
    ids.insert(ids.end(), local.src_image_ids.begin(), local.src_image_ids.end());
    for (size_t i = 0; i < ids.size(); ++i) {
        // a source view is limited by ITS OWN problem's current size (ACMMP.cpp:569-571: indexed by image id)
        const int max_size = i == 0 ? local.cur_image_size : problem[ids[i]].cur_image_size;
        cv::Mat image;
        Camera camera;
        load_view(dense_folder, ids[i], max_size, image, camera);
        images.push_back(image);
        cameras.push_back(camera);
    }
    params.depth_min = cameras[0].depth_min * 0.6f;
    params.depth_max = cameras[0].depth_max * 1.2f;
    params.num_images = (int)images.size();
    params.disparity_min = cameras[0].K[0] * params.baseline / params.depth_max;
    params.disparity_max = cameras[0].K[0] * params.baseline / params.depth_min;
    if (params.geom_consistency) {   // depth maps of the reference view and its sources from the previous stage
        depths.clear();
        const char *suffix = params.multi_geometry ? "/depths_geom.dmb" : "/depths.dmb";
        for (int id : ids) {
            cv::Mat_<float> depth;
            readDepthDmb(result_dir(output_folder, id) + suffix, depth);
            depths.push_back(depth);
        }
    }
}

// uploads + per-pixel state (ACMMP.cpp:638-809): textures, depth maps, warm start / hierarchy inputs from the .dmb files
void ACMMP::CudaSpaceInitialization(const std::string &output_folder, const Problem &problem) {
    acmmp_ctx *ctx = ctx_of(this);
    num_images = (int)images.size();
    std::vector<const float *> ptrs;
    for (const cv::Mat &m : images) ptrs.push_back(m.ptr<float>());
    CK(ctx, acmmp_set_problem(ctx, num_images, reinterpret_cast<const acmmp_camera *>(cameras.data()), ptrs.data(), 0));
    const size_t P = (size_t)cameras[0].width * cameras[0].height;
    plane_hypotheses_host = new float4[P];
    costs_host = new float[P];
    const std::string folder = result_dir(output_folder, problem.ref_image_id);
    if (params.geom_consistency) {
        for (int i = 0; i < num_images; ++i) CK(ctx, acmmp_set_depth_map(ctx, i, depths[i].ptr<float>(), depths[i].cols, depths[i].rows, 0));
        cv::Mat_<float> ref_depth, ref_cost;
        cv::Mat_<cv::Vec3f> ref_normal;
        readDepthDmb(folder + (params.multi_geometry ? "/depths_geom.dmb" : "/depths.dmb"), ref_depth);
        readNormalDmb(folder + "/normals.dmb", ref_normal);
        readDepthDmb(folder + "/costs.dmb", ref_cost);
        CK(ctx, acmmp_set_warm_start(ctx, ref_depth.ptr<float>(), ref_normal.ptr<float>(), ref_cost.ptr<float>()));
    }
    if (params.hierarchy) {
        // depths.dmb is already at this scale (written by RunJBU); normals / costs are the previous scale's
        cv::Mat_<float> ref_depth, ref_cost;
        cv::Mat_<cv::Vec3f> ref_normal;
        readDepthDmb(folder + "/depths.dmb", ref_depth);
        readNormalDmb(folder + "/normals.dmb", ref_normal);
        readDepthDmb(folder + "/costs.dmb", ref_cost);
        const int width = ref_normal.cols, height = ref_normal.rows;
        if (width != images[0].rows || height != images[0].cols) {   // sic (ACMMP.cpp:768): true for every non-square image
            params.upsample = true;
            params.scaled_cols = width;
            params.scaled_rows = height;
        } else {
            params.upsample = false;
        }
        CK(ctx, acmmp_set_hierarchy_inputs(ctx, ref_normal.ptr<float>(), ref_cost.ptr<float>(), width, height, ref_depth.ptr<float>()));
    }
}

void ACMMP::RunPatchMatch() {   // ACMMP.cu:1378-1456
    acmmp_ctx *ctx = ctx_of(this);
    CK(ctx, acmmp_run(ctx, g_seed * 1000003ull + (g_runs++)));
    CK(ctx, acmmp_get_result(ctx, reinterpret_cast<float *>(plane_hypotheses_host), costs_host, nullptr));
}

int ACMMP::GetReferenceImageWidth() { return cameras[0].width; }
int ACMMP::GetReferenceImageHeight() { return cameras[0].height; }
cv::Mat ACMMP::GetReferenceImage() { return images[0]; }
float4 ACMMP::GetPlaneHypothesis(const int index) { return plane_hypotheses_host[index]; }
Camera ACMMP::GetCamera(const int index) { return cameras[index]; }
float ACMMP::GetCost(const int index) { return costs_host[index]; }
float ACMMP::GetMinDepth() { return params.depth_min; }
float ACMMP::GetMaxDepth() { return params.depth_max; }

// best pixel (cost < 0.1) of every 5 x 5 cell, cells in column-major order (ACMMP.cpp:868-894)
void ACMMP::GetSupportPoints(std::vector<cv::Point> &support2DPoints) {
    support2DPoints.clear();
    const int step = 5, width = cameras[0].width, height = cameras[0].height;
    for (int col = 0; col < width; col += step)
        for (int row = 0; row < height; row += step) {
            float best = 2.0f;
            cv::Point where;
            for (int c = col; c < std::min(width, col + step); ++c)
                for (int r = row; r < std::min(height, row + step); ++r) {
                    const float cost = costs_host[r * width + c];
                    if (cost < 2.0f && best > cost) {
                        where = cv::Point(c, r);
                        best = cost;
                    }
                }
            if (best < 0.1f) support2DPoints.push_back(where);
        }
}

// cv::Subdiv2D of the reference (ACMMP.cpp:896-918) -> the library's integer Delaunay (acmmp_delaunay)
std::vector<Triangle> ACMMP::DelaunayTriangulation(const cv::Rect boundRC, const std::vector<cv::Point> &points) {
    (void)boundRC;
    std::vector<Triangle> results;
    if (points.empty()) return results;
    std::vector<int32_t> xy(points.size() * 2), tris(points.size() * 12 + 12);
    for (size_t i = 0; i < points.size(); ++i) {
        xy[2 * i] = points[i].x;
        xy[2 * i + 1] = points[i].y;
    }
    const int nt = acmmp_delaunay(xy.data(), (int)points.size(), tris.data(), (int)tris.size() / 6);
    if (nt < 0) die("acmmp_delaunay", nullptr);
    for (int t = 0; t < nt && t < (int)tris.size() / 6; ++t)
        results.push_back(Triangle(cv::Point(tris[6 * t], tris[6 * t + 1]), cv::Point(tris[6 * t + 2], tris[6 * t + 3]), cv::Point(tris[6 * t + 4], tris[6 * t + 5])));
    return results;
}

// plane through the three back-projected vertices, n.X + d = 0 with d > 0 (ACMMP.cpp:920-953): the null vector of
// [X 1] is the generalised cross product of its three rows
float4 ACMMP::GetPriorPlaneParams(const Triangle triangle, const cv::Mat_<float> depths) {
    const cv::Point pts[3] = {triangle.pt1, triangle.pt2, triangle.pt3};
    double a[3][4];
    for (int r = 0; r < 3; ++r) {
        const float depth = depths(pts[r].y, pts[r].x);
        a[r][0] = depth * (pts[r].x - cameras[0].K[2]) / cameras[0].K[0];
        a[r][1] = depth * (pts[r].y - cameras[0].K[5]) / cameras[0].K[4];
        a[r][2] = depth;
        a[r][3] = 1.0f;
    }
    auto det3 = [&](int c0, int c1, int c2) {
        return a[0][c0] * (a[1][c1] * a[2][c2] - a[1][c2] * a[2][c1]) - a[0][c1] * (a[1][c0] * a[2][c2] - a[1][c2] * a[2][c0]) +
               a[0][c2] * (a[1][c0] * a[2][c1] - a[1][c1] * a[2][c0]);
    };
    double n[4] = {det3(1, 2, 3), -det3(0, 2, 3), det3(0, 1, 3), -det3(0, 1, 2)};
    double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2] + n[3] * n[3]);
    if (len == 0) len = 1;
    float4 n4 = make_float4((float)(n[0] / len), (float)(n[1] / len), (float)(n[2] / len), (float)(n[3] / len));
    float norm2 = sqrt(pow(n4.x, 2) + pow(n4.y, 2) + pow(n4.z, 2));
    if (n4.w < 0) norm2 *= -1;
    n4.x /= norm2;
    n4.y /= norm2;
    n4.z /= norm2;
    n4.w /= norm2;
    return n4;
}

float ACMMP::GetDepthFromPlaneParam(const float4 plane_hypothesis, const int x, const int y) {   // ACMMP.cpp:955-958
    return -plane_hypothesis.w * cameras[0].K[0] /
           ((x - cameras[0].K[2]) * plane_hypothesis.x + (cameras[0].K[0] / cameras[0].K[4]) * (y - cameras[0].K[5]) * plane_hypothesis.y +
            cameras[0].K[0] * plane_hypothesis.z);
}

// per-pixel prior planes + triangle labels -> the device (ACMMP.cpp:811-834); uncovered pixels keep a zero plane
void ACMMP::CudaPlanarPriorInitialization(const std::vector<float4> &PlaneParams, const cv::Mat_<float> &masks) {
    const int width = cameras[0].width, height = cameras[0].height;
    prior_planes_host = new float4[(size_t)width * height]();
    plane_masks_host = new unsigned int[(size_t)width * height];
    for (int j = 0; j < height; ++j)
        for (int i = 0; i < width; ++i) {
            const int center = j * width + i;
            plane_masks_host[center] = (unsigned int)masks(j, i);
            if (masks(j, i) > 0) prior_planes_host[center] = PlaneParams[masks(j, i) - 1];
        }
    CK(ctx_of(this), acmmp_set_planar_prior(ctx_of(this), reinterpret_cast<const float *>(prior_planes_host), plane_masks_host));
}

void ACMMP::Colmap2MVS(const std::string &, std::vector<Problem> &) {}   // declared, never defined by the reference either
