// acmmp_main.cpp - command line of the B200 path with the reference's flags and file layout
// (main_ACMMP.cpp:9-199): reads <dense_folder>/pair.txt, cams/%08d_cam.txt, images/%08d.*,
// runs the multi-scale schedule (photometric + planar prior, geom x2 per scale) resident on
// the device, writes <dense_folder><output_dir>/2333_%08d/{depths,depths_geom,normals,costs}.dmb
// and fuses into ACMMP_model.ply.  Hand-rolled flag parser (Boost is not in this image).
// -p / --prior: seeded mode - <dense_folder>/priors/{depths,normals}/%08d.png (16 bit) give the initial
// hypotheses of the first scale (pSampler, acmmp_definitions.cpp:7-177); output goes to /ACMMP_PRIOR
// unless --output_dir is given (main_ACMMP.cpp:79-85).
// --force_fusion / --multi_fusion [DIR] / --single_match_penalty N select the prior-aware fusion
// (main_ACMMP.cpp:178-189): map set 0 is read from <dense_folder><DIR>/2333_*/, set 1 is this run's.
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <thread>
#include <string>
#include <vector>

#include "fuse_common.hpp"
#include "../acmmp_b200/host/driver.hpp"

using namespace acmmp_b200;
using namespace acmmp_cli;

// --reference_flow: main() of the reference as it stands (main_ACMMP.cpp:89-176) - one ProcessProblem per view and stage,
// .dmb files between stages, JointBilateralUpsampling between scales - through `class ACMMP` (host/acmmp.hpp, host/driver.hpp).
// The default mode runs the same schedule resident on the device (class Scan).
static int reference_flow(const std::string &dense_folder, const std::string &output_folder, std::vector<Problem> &problems, uint64_t seed, int device) {
    const size_t num_images = problems.size();
    int max_num_downscale = ComputeMultiScaleSettings(dense_folder, problems, device);
    bool first_scale = true;
    uint64_t run = 0;
    while (max_num_downscale >= 0) {
        printf("Scale: %d\n", max_num_downscale);
        for (size_t i = 0; i < num_images; ++i)
            if (problems[i].num_downscale >= 0) {
                problems[i].cur_image_size = problems[i].max_image_size / pow(2, problems[i].num_downscale);
                problems[i].num_downscale--;
            }
        bool hierarchy = false;
        if (!first_scale) {
            for (size_t i = 0; i < num_images; ++i) JointBilateralUpsampling(dense_folder, output_folder, problems[i], problems[i].cur_image_size, device);
            hierarchy = true;
        }
        first_scale = false;
        for (size_t i = 0; i < num_images; ++i)
            ProcessProblem(output_folder, dense_folder, problems, (int)i, false, true, hierarchy, false, seed * 1000003ull + 2 * (run++), true, device);
        for (int geom_iter = 0; geom_iter < 2; ++geom_iter)
            for (size_t i = 0; i < num_images; ++i)
                ProcessProblem(output_folder, dense_folder, problems, (int)i, true, false, false, geom_iter != 0, seed * 1000003ull + 2 * (run++), true, device);
        max_num_downscale--;
    }
    return 0;
}

int main(int argc, char **argv) {
    std::string dense_folder, output_dir = "/ACMMP", mask_dir = " ", image_dir = "/images";
    float consistency_scalar = 0.3f;
    int num_consistent_thresh = 1, single_match_penalty = 0, device = 0, gpus = 1;
    unsigned long long seed = 0;
    bool fuse = true, force_fusion = false, multi_fusion = false, prior = false, ref_flow = false, staged = false;
    std::string fusion_dir = "/ACMMP";
    bool renamed_outdir = false;
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
            printf("usage: acmmp_main <dense_folder> [--output_dir /ACMMP] [-f|--fuse_thresh 0.3] [--num_consistent_thresh 1]\n"
                   "                  [--mask_dir DIR] [--image_override DIR] [--device 0] [--gpus N] [--seed N] [--staged] [--no_fusion] [--reference_flow]\n"
                   "                  [--force_fusion] [--multi_fusion [/DIR]] [--single_match_penalty N] [-p|--prior]\n");
            return 1;
        } else if (a == "--dense_folder") dense_folder = value("--dense_folder");
        else if (a == "--output_dir") {
            output_dir = value("--output_dir");
            renamed_outdir = true;
        }
        else if (a == "-f" || a == "--fuse_thresh") consistency_scalar = (float)atof(value("--fuse_thresh"));
        else if (a == "--num_consistent_thresh") num_consistent_thresh = atoi(value("--num_consistent_thresh"));
        else if (a == "--mask_dir") mask_dir = value("--mask_dir");
        else if (a == "--image_override") image_dir = value("--image_override");
        else if (a == "--device") device = atoi(value("--device"));
        else if (a == "--gpus") gpus = atoi(value("--gpus"));
        else if (a == "--staged") staged = true;
        else if (a == "--seed") seed = strtoull(value("--seed"), nullptr, 10);
        else if (a == "--no_fusion") fuse = false;
        else if (a == "--reference_flow") ref_flow = true;
        else if (a == "--force_fusion") force_fusion = true;
        else if (a == "--multi_fusion") {   // implicit value "/ACMMP" (main_ACMMP.cpp:31)
            multi_fusion = true;
            if (i + 1 < argc && argv[i + 1][0] == '/') fusion_dir = argv[++i];
        } else if (a == "--single_match_penalty") single_match_penalty = atoi(value("--single_match_penalty"));
        else if (a == "-p" || a == "--prior") prior = true;
        else if (a[0] != '-') dense_folder = a;
        else {
            fprintf(stderr, "unknown option %s\n", a.c_str());
            return EXIT_FAILURE;
        }
    }
    if (dense_folder.empty()) {
        fprintf(stderr, "no dense_folder given\n");
        return EXIT_FAILURE;
    }
    if (gpus < 1) {
        fprintf(stderr, "--gpus must be >= 1\n");
        return EXIT_FAILURE;
    }
    try {
        std::vector<Problem> problems;
        if (!GenerateSampleList(dense_folder, problems) || problems.empty()) {
            fprintf(stderr, "cannot read %s/pair.txt\n", dense_folder.c_str());
            return EXIT_FAILURE;
        }
        const int n = (int)problems.size();
        printf("There are %d problems needed to be processed!\n", n);
        std::map<int, int> image_id_2_index;
        for (int i = 0; i < n; ++i) image_id_2_index[problems[i].ref_image_id] = i;
        std::vector<acmmp_camera> cams(n);
        std::vector<Map<uint8_t>> images(n);
        std::vector<std::vector<int>> pairs(n);
        for (int i = 0; i < n; ++i) {
            const int id = problems[i].ref_image_id;
            if (!ReadCamera(dense_folder + "/cams/" + id_name(id, "_cam.txt"), cams[i])) {
                fprintf(stderr, "cannot read camera of image %d\n", id);
                return EXIT_FAILURE;
            }
            std::string err;
            // grey read for PatchMatch (ACMMP.cpp:534-541); a JPEG is decoded straight to luma like imread(IMREAD_GRAYSCALE)
            if (!find_image(dense_folder + "/images", id, images[i], err, false, device)) {
                fprintf(stderr, "%s\n", err.c_str());
                return EXIT_FAILURE;
            }
            cams[i].width = images[i].cols;
            cams[i].height = images[i].rows;
            for (int sid : problems[i].src_image_ids) {
                auto it = image_id_2_index.find(sid);
                if (it == image_id_2_index.end()) {
                    fprintf(stderr, "pair.txt: source %d of image %d is not a reference image\n", sid, id);
                    return EXIT_FAILURE;
                }
                pairs[i].push_back(it->second);
            }
        }
        if (prior && !renamed_outdir) output_dir = "/ACMMP_PRIOR";   // main_ACMMP.cpp:79-85
        const std::string output_folder = dense_folder + output_dir;
        mkdir(output_folder.c_str(), 0777);

        if (ref_flow) {
            if (prior || force_fusion || multi_fusion) {
                fprintf(stderr, "--reference_flow runs the plain schedule (no -p / --force_fusion / --multi_fusion)\n");
                return EXIT_FAILURE;
            }
            reference_flow(dense_folder, output_folder, problems, seed, device);
            printf("depth maps written to %s\n", output_folder.c_str());
            if (fuse) {
                std::vector<acmmp_camera> cams_out(n);
                std::vector<Map<float>> depths(n), normals(n);
                for (int i = 0; i < n; ++i) {
                    const std::string folder = ACMMP::ResultFolder(output_folder, problems[i].ref_image_id);
                    if (readDepthDmb(folder + "/depths_geom.dmb", depths[i]) || readNormalDmb(folder + "/normals.dmb", normals[i])) {
                        fprintf(stderr, "cannot read the maps under %s\n", folder.c_str());
                        return EXIT_FAILURE;
                    }
                    cams_out[i] = CameraAtMapSize(cams[i], depths[i].cols, depths[i].rows);
                }
                FuseOptions o;
                o.dense_folder = dense_folder;
                o.output_folder = output_folder;
                o.image_dir = image_dir;
                o.mask_dir = mask_dir;
                o.consistency_scalar = consistency_scalar;
                o.num_consistent_thresh = num_consistent_thresh;
                o.device = device;
                if (FuseAndStore(o, problems, cams_out, pairs, [&](int i, Map<float> &d, Map<float> &nrm) {
                        d = depths[i];
                        nrm = normals[i];
                        return true;
                    }) < 0)
                    return EXIT_FAILURE;
            }
            return 0;
        }
        // --gpus N: one host thread and one resident scan per device, communicators from ncclCommInitAll.  The scan runs
        // as a dataflow of (stage, view) tasks (Scan::RunDataflow: rounds of N tasks across stage boundaries, the records
        // of a round broadcast by the library while the next round computes).  --staged keeps the stage-by-stage loop
        // (view v belongs to rank v % N, ncclAllGather of the depth maps after every stage); N = 1 always runs that
        // loop, in the reference's in-place order, without a communicator.
        std::vector<std::unique_ptr<Scan>> scans;
        for (int r = 0; r < gpus; ++r) {
            scans.emplace_back(new Scan(cams, pairs, gpus > 1 ? r : device, r, gpus));
            for (int i = 0; i < n; ++i) scans[r]->SetImageU8(i, ToGray8(images[i]));
        }
        if (gpus > 1) Scan::CommInitAll(scans);
        if (prior) {
            // the reference numbers the prior images by problem index, not by image id (acmmp_definitions.cpp:278, :104-105)
            for (int i = 0; i < n; ++i) {
                Map<uint16_t> d16, n16;
                std::string err;
                if (!ReadPng16(dense_folder + "/priors/depths/" + id_name(i, ".png"), d16, nullptr, &err) ||
                    !ReadPng16(dense_folder + "/priors/normals/" + id_name(i, ".png"), n16, nullptr, &err)) {
                    fprintf(stderr, "Initialisation from a prior was requested, but no suitable priors were found: %s\n", err.c_str());
                    return -1;
                }
                for (auto &sc : scans) sc->SetPriorMaps(i, d16, n16);
            }
        }
        std::vector<Map<float>> photo_depth(n);
        const int levels = scans[0]->NumLevels();
        std::vector<std::string> errors(gpus);
        const bool dataflow = gpus > 1 && !staged;
        auto drive = [&](int r) {
            try {
                Scan &scan = *scans[r];
                if (dataflow) {
                    if (r == 0) scan.KeepPhotometric(true);
                    scan.RunDataflow(seed);
                    if (r == 0)
                        for (int i = 0; i < n; ++i) scan.GetPhotometricDepth(i, photo_depth[i]);
                    scan.Sync();
                    return;
                }
                for (int level = 0; level < levels; ++level) {
                    if (r == 0) printf("Scale: %d\n", levels - 1 - level);
                    scan.BeginLevel(level);
                    for (int stage = ACMMP_STAGE_PHOTO; stage <= ACMMP_STAGE_GEOM1; ++stage) {
                        scan.RunStage(stage, seed);
                        scan.AllGather();
                        if (r == 0 && level == levels - 1 && stage == ACMMP_STAGE_PHOTO)   // depths.dmb = the photometric + planar-prior result
                            for (int i = 0; i < n; ++i) scan.GetView(i, &photo_depth[i], nullptr, nullptr);
                    }
                }
                scan.Sync();
            } catch (const std::exception &e) {
                errors[r] = e.what();
            }
        };
        if (gpus == 1) {
            drive(0);
        } else {
            std::vector<std::thread> threads;
            for (int r = 0; r < gpus; ++r) threads.emplace_back(drive, r);
            for (auto &t : threads) t.join();
        }
        for (int r = 0; r < gpus; ++r)
            if (!errors[r].empty()) throw std::runtime_error("rank " + std::to_string(r) + ": " + errors[r]);
        // every rank holds every depth map (exchange buffer); normals and costs live on the owner
        // (after a dataflow run every rank holds every view's maps)
        auto get_view = [&](int i, Map<float> *d, Map<float> *nrm, Map<float> *c) { scans[dataflow ? 0 : i % gpus]->GetView(i, d, nrm, c); };
        // acmmp_definitions.cpp:392-401: one folder per reference image
        std::vector<acmmp_camera> cams_out(n);
        for (int i = 0; i < n; ++i) {
            const std::string folder = output_folder + "/2333_" + id_name(problems[i].ref_image_id, "");
            mkdir(folder.c_str(), 0777);
            Map<float> depth, normal, cost;
            get_view(i, &depth, &normal, &cost);
            if (writeDepthDmb(folder + "/depths.dmb", photo_depth[i]) || writeDepthDmb(folder + "/depths_geom.dmb", depth) ||
                writeNormalDmb(folder + "/normals.dmb", normal) || writeDepthDmb(folder + "/costs.dmb", cost)) {
                fprintf(stderr, "cannot write maps under %s\n", folder.c_str());
                return EXIT_FAILURE;
            }
            cams_out[i] = CameraAtMapSize(cams[i], depth.cols, depth.rows);
        }
        printf("depth maps written to %s\n", output_folder.c_str());
        if (fuse) {
            FuseOptions o;
            o.dense_folder = dense_folder;
            o.output_folder = output_folder;
            o.fusion_folder = dense_folder + fusion_dir;
            o.image_dir = image_dir;
            o.mask_dir = mask_dir;
            o.prior_aware = (prior && multi_fusion) || force_fusion;   // main_ACMMP.cpp:181
            o.consistency_scalar = consistency_scalar;
            o.num_consistent_thresh = num_consistent_thresh;
            o.single_match_penalty = single_match_penalty;
            o.device = device;
            // colour and mask come from the host, so the maps take the same way (one flag per call)
            if (FuseAndStore(o, problems, cams_out, pairs, [&](int i, Map<float> &d, Map<float> &nrm) {
                    get_view(i, &d, &nrm, nullptr);
                    return true;
                }) < 0)
                return EXIT_FAILURE;
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "%s\n", e.what());   // the reference: CudaSafeCall prints and exit(EXIT_FAILURE)
        return EXIT_FAILURE;
    }
    return 0;
}
