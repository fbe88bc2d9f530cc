// fuse_data.cpp - the fusion-only command line of the reference (main_fusion.cpp:7-95) over the C ABI: re-fuses an
// existing output folder.  Same flags; same conventions:
//   * the maps are <dense_folder>/ACMMP/2333_%08d/depths.dmb + normals.dmb - main_fusion.cpp calls the fusion with
//     geom_consistency = false (:75), i.e. the PHOTOMETRIC depth maps, and has no --output_dir (:11,:66-72);
//   * -f / --fuse_thresh, --num_consistent_thresh, --single_match_penalty, --mask_dir, --image_override;
//   * --force_fusion (with --multi_fusion [/DIR] naming the second folder): prior-aware fusion of the maps under
//     <dense_folder>/DIR (set 0) and under /ACMMP (set 1), written to ACMMP_prior_model.ply.  The reference passes
//     `image_override` where RunPriorAwareFusion expects the mask folder (main_fusion.cpp:79-84), so its run reads
//     masks from "/images/%08d.png" and indexes an empty cv::Mat; here that mode runs without masks.
// --device N is an addition.
#include <sys/stat.h>

#include <cstdlib>
#include <map>

#include "fuse_common.hpp"

using namespace acmmp_b200;
using namespace acmmp_cli;

int main(int argc, char **argv) {
    FuseOptions o;
    std::string fusion_dir = "/ACMMP";
    bool force_fusion = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto value = [&](const char *name) -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "missing value for %s\n", name);
                exit(EXIT_FAILURE);
            }
            return argv[++i];
        };
        if (a == "-h" || a == "--help") {
            printf("usage: fuse_data <dense_folder> [-f|--fuse_thresh 0.3] [--num_consistent_thresh 1] [--single_match_penalty 0]\n"
                   "                 [--mask_dir DIR] [--image_override DIR] [--force_fusion] [--multi_fusion [/DIR]] [--device 0]\n");
            return 1;
        } else if (a == "--dense_folder") o.dense_folder = value("--dense_folder");
        else if (a == "-f" || a == "--fuse_thresh") o.consistency_scalar = (float)atof(value("--fuse_thresh"));
        else if (a == "--num_consistent_thresh") o.num_consistent_thresh = atoi(value("--num_consistent_thresh"));
        else if (a == "--single_match_penalty") o.single_match_penalty = atoi(value("--single_match_penalty"));
        else if (a == "--mask_dir") o.mask_dir = value("--mask_dir");
        else if (a == "--image_override") o.image_dir = value("--image_override");
        else if (a == "--device") o.device = atoi(value("--device"));
        else if (a == "--force_fusion") force_fusion = true;
        else if (a == "--multi_fusion") {   // implicit value "/ACMMP" (main_fusion.cpp:29)
            if (i + 1 < argc && argv[i + 1][0] == '/') fusion_dir = argv[++i];
        } else if (a[0] != '-') o.dense_folder = a;
        else {
            fprintf(stderr, "unknown option %s\n", a.c_str());
            return EXIT_FAILURE;
        }
    }
    if (o.dense_folder.empty()) {
        fprintf(stderr, "no dense_folder given\n");
        return EXIT_FAILURE;
    }
    try {
        std::vector<Problem> problems;
        if (!GenerateSampleList(o.dense_folder, problems) || problems.empty()) {
            fprintf(stderr, "cannot read %s/pair.txt\n", o.dense_folder.c_str());
            return EXIT_FAILURE;
        }
        const int n = (int)problems.size();
        printf("There are %d problems needed to be processed!\n", n);
        o.output_folder = o.dense_folder + "/ACMMP";
        o.fusion_folder = o.dense_folder + fusion_dir;
        o.prior_aware = force_fusion;              // `prior` is never set by main_fusion.cpp (:58): only --force_fusion selects it
        o.set0_depth_name = "/depths.dmb";          // geom_consistency = false (:75)
        if (o.prior_aware) o.mask_dir = " ";
        mkdir(o.output_folder.c_str(), 0777);
        std::map<int, int> image_id_2_index;
        for (int i = 0; i < n; ++i) image_id_2_index[problems[i].ref_image_id] = i;
        std::vector<acmmp_camera> cams_out(n);
        std::vector<std::vector<int>> pairs(n);
        std::vector<Map<float>> depths(n), normals(n);
        for (int i = 0; i < n; ++i) {
            const int id = problems[i].ref_image_id;
            acmmp_camera cam;
            if (!ReadCamera(o.dense_folder + "/cams/" + id_name(id, "_cam.txt"), cam)) {
                fprintf(stderr, "cannot read camera of image %d\n", id);
                return EXIT_FAILURE;
            }
            const std::string folder = o.output_folder + "/2333_" + id_name(id, "");
            if (readDepthDmb(folder + "/depths.dmb", depths[i]) || readNormalDmb(folder + "/normals.dmb", normals[i])) {
                fprintf(stderr, "cannot read the maps of image %d under %s\n", id, folder.c_str());
                return EXIT_FAILURE;
            }
            // the camera file carries no size: it is the colour image's (acmmp_definitions.cpp:857-878)
            Map<uint8_t> color;
            std::string err;
            if (!find_image(o.dense_folder + o.image_dir, id, color, err, true, o.device)) {
                fprintf(stderr, "%s\n", err.c_str());
                return EXIT_FAILURE;
            }
            cam.width = color.cols;
            cam.height = color.rows;
            cams_out[i] = CameraAtMapSize(cam, depths[i].cols, depths[i].rows);
            for (int sid : problems[i].src_image_ids) {
                auto it = image_id_2_index.find(sid);
                if (it == image_id_2_index.end()) {
                    fprintf(stderr, "pair.txt: source %d of image %d is not a reference image\n", sid, id);
                    return EXIT_FAILURE;
                }
                pairs[i].push_back(it->second);
            }
        }
        if (FuseAndStore(o, problems, cams_out, pairs, [&](int i, Map<float> &d, Map<float> &nrm) {
                d = depths[i];
                nrm = normals[i];
                return true;
            }) < 0)
            return EXIT_FAILURE;
    } catch (const std::exception &e) {
        fprintf(stderr, "%s\n", e.what());
        return EXIT_FAILURE;
    }
    return 0;
}
