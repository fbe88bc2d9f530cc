// io_selftest.cpp - exercises acmmp_b200/host/io.hpp from the CPU test-suite (tests/test_host_io.py):
//   io_selftest <dir>   reads <dir>/pair.txt, <dir>/cams/00000000_cam.txt, <dir>/in.dmb, <dir>/img.ppm, <dir>/points.bin
//   io_selftest --png in.png out.bin   decodes a PNG (ReadPng16) and dumps the samples
//   io_selftest --resize8 in.ppm out.bin W H   cv::resize(INTER_LINEAR) of an 8-bit image (ResizeLinear8)
//   io_selftest --dmb file depth|normal|any    header validation of the .dmb readers
//                       writes <dir>/out.dmb, <dir>/gray.dmb, <dir>/out.ply and prints what it parsed as key=value lines
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#define ACMMP_IO_WITH_PNG 1
#include "../acmmp_b200/host/io.hpp"

using namespace acmmp_b200;

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    if (argc >= 4 && std::string(argv[1]) == "--png") {   // io_selftest --png in.png out.bin: raw samples (u16, cv::imread(-1) order)
        Map<uint16_t> png;
        int depth = 0;
        std::string perr;
        if (!ReadPng16(argv[2], png, &depth, &perr)) {
            fprintf(stderr, "%s\n", perr.c_str());
            return 12;
        }
        printf("png=%d,%d,%d,%d\n", png.rows, png.cols, png.channels, depth);
        FILE *o = fopen(argv[3], "wb");
        if (!o) return 13;
        fwrite(png.data.data(), sizeof(uint16_t), png.data.size(), o);
        fclose(o);
        return 0;
    }
    if (argc >= 6 && std::string(argv[1]) == "--resize8") {   // io_selftest --resize8 in.ppm|pgm out.bin W H: ResizeLinear8
        Map<uint8_t> im;
        std::string err;
        if (!ReadImage(argv[2], im, &err)) {
            fprintf(stderr, "%s\n", err.c_str());
            return 14;
        }
        const Map<uint8_t> out = ResizeLinear8(im, atoi(argv[4]), atoi(argv[5]));
        FILE *o = fopen(argv[3], "wb");
        if (!o) return 13;
        fwrite(out.data.data(), 1, out.data.size(), o);
        fclose(o);
        printf("resized=%d,%d,%d\n", out.rows, out.cols, out.channels);
        return 0;
    }
    if (argc >= 4 && std::string(argv[1]) == "--cache") {   // io_selftest --cache a.pgm b.pgm: the decoded-image content cache
        ImageCache &cache = ImageCache::Instance();
        Map<uint8_t> a1, a2, b1, a3;
        std::string err;
        if (!ReadImageCached(argv[2], a1, &err, false) || !ReadImageCached(argv[2], a2, &err, false) || !ReadImageCached(argv[3], b1, &err, false)) {
            fprintf(stderr, "%s\n", err.c_str());
            return 16;
        }
        ImageCache::Stats st = cache.stats();
        printf("after_three_reads=%llu,%llu,%zu same=%d\n", (unsigned long long)st.hits, (unsigned long long)st.misses, st.bytes, a1.data == a2.data ? 1 : 0);
        // rewrite the first file with other content (and another size): the entry must not be served again
        FILE *f = fopen(argv[2], "wb");
        if (!f) return 13;
        fprintf(f, "P5\n2 1\n255\n");
        fputc(7, f);
        fputc(9, f);
        fclose(f);
        if (!ReadImageCached(argv[2], a3, &err, false)) return 16;
        st = cache.stats();
        printf("after_rewrite=%llu,%llu,%zu first=%d,%d,%d\n", (unsigned long long)st.hits, (unsigned long long)st.misses, st.bytes, a3.rows, a3.cols, (int)a3.data[0]);
        cache.SetBudget(b1.data.size());   // room for one image: the least recently used one goes
        st = cache.stats();
        printf("after_budget=%llu,%zu\n", (unsigned long long)st.evictions, st.bytes);
        Map<uint8_t> b2;
        if (!ReadImageCached(argv[3], b2, &err, false)) return 16;
        st = cache.stats();
        printf("final=%llu,%llu\n", (unsigned long long)st.hits, (unsigned long long)st.misses);
        return 0;
    }
    if (argc >= 4 && std::string(argv[1]) == "--dmb") {   // io_selftest --dmb file depth|normal|any: 0 if accepted
        Map<float> m;
        const std::string kind = argv[3];
        const int rc = kind == "depth" ? readDepthDmb(argv[2], m) : kind == "normal" ? readNormalDmb(argv[2], m) : readDmb(argv[2], m);
        printf("dmb_rc=%d\n", rc);
        return rc ? 15 : 0;
    }
    const std::string dir = argv[1];
    std::vector<Problem> problems;
    if (!GenerateSampleList(dir, problems)) return 3;
    printf("problems=%zu\n", problems.size());
    for (const auto &p : problems) {
        printf("ref=%d src=", p.ref_image_id);
        for (int s : p.src_image_ids) printf("%d,", s);
        printf("\n");
    }
    acmmp_camera cam;
    if (!ReadCamera(dir + "/cams/00000000_cam.txt", cam)) return 4;
    printf("K=");
    for (float v : cam.K) printf("%.9g,", v);
    printf("\nR=");
    for (float v : cam.R) printf("%.9g,", v);
    printf("\nt=%.9g,%.9g,%.9g\ndepth=%.9g,%.9g\n", cam.t[0], cam.t[1], cam.t[2], cam.depth_min, cam.depth_max);
    Map<float> m;
    if (readDmb(dir + "/in.dmb", m)) return 5;
    printf("dmb=%d,%d,%d\n", m.rows, m.cols, m.channels);
    if (writeDmb(dir + "/out.dmb", m.data.data(), m.rows, m.cols, m.channels)) return 6;
    Map<uint8_t> im;
    std::string err;
    if (!ReadImage(dir + "/img.ppm", im, &err)) {
        fprintf(stderr, "%s\n", err.c_str());
        return 7;
    }
    printf("image=%d,%d,%d\n", im.rows, im.cols, im.channels);
    if (writeDepthDmb(dir + "/gray.dmb", ToGrayFloat(im))) return 8;
    FILE *f = fopen((dir + "/points.bin").c_str(), "rb");
    if (!f) return 9;
    std::vector<float> pts;
    float rec[9];
    while (fread(rec, sizeof(float), 9, f) == 9) pts.insert(pts.end(), rec, rec + 9);
    fclose(f);
    if (!StoreColorPlyFileBinaryPointCloud(dir + "/out.ply", pts.data(), pts.size() / 9)) return 10;
    printf("points=%zu\n", pts.size() / 9);
    if (ReadImage(dir + "/pair.txt", im, &err)) return 11;   // must be rejected
    return 0;
}
