// vision-cli: headless stand-in for the reference's GUI-only driver (src/apps/vision-gui/main.cpp:10-14,
// application.cpp:37-49,174-199; CLI flags of src/base/cli_parser.cpp:18-43 that this path uses).
//   vision-cli -s <scene.json> [-o out.ppm] [--spp N] [-d cuda] [--device-index I] [--pfm out.pfm] [--gpus N [--nccl]] [--two-level] [--opt key=value]
// --two-level: one BVH per unique mesh + a top-level BVH over the instances (the reference's GAS / IAS shape) instead of the flattened default;
// --opt: any vb_ctx_set_option pair, applied before the BVH is built.
// --gpus N: the spp frame indices are split over N GPUs of this box (rank g renders g, g + N, ...), films reduced + resolved over
// NVLink peer memory (vb_film_reduce_peers) or, with --nccl, by ncclAllReduce (vb_film_allreduce).
#include "vh_scene.h"
#include <chrono>
#include <cstdio>
#include <cstring>

int main(int argc, char **argv) {
    std::string scene, out = "output.ppm", pfm, aov, device = "cuda";
    uint32_t spp = 0, gpus = 1; int dev_index = 0; bool nccl = false;
    std::vector<std::pair<std::string, double>> options;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str()); exit(2); } return std::string(argv[++i]); };
        if (a == "-s" || a == "--scene") scene = next();
        else if (a == "-o" || a == "--output") out = next();
        else if (a == "--pfm") pfm = next();
        else if (a == "--aov") aov = next();
        else if (a == "--spp") spp = (uint32_t)std::stoul(next());
        else if (a == "-d" || a == "--device") device = next();
        else if (a == "--device-index") dev_index = std::stoi(next());
        else if (a == "--gpus") gpus = (uint32_t)std::stoul(next());
        else if (a == "--nccl") nccl = true;
        else if (a == "--two-level") options.emplace_back("build.two_level", 1.0);
        else if (a == "--opt") { std::string kv = next(); size_t eq = kv.find('='); if (eq == std::string::npos) { fprintf(stderr, "--opt wants key=value\n"); return 2; } options.emplace_back(kv.substr(0, eq), std::stod(kv.substr(eq + 1))); }
        else if (a == "-h" || a == "--help") { printf("vision-cli -s scene.json [-o out.png|.ppm|.hdr|.pfm] [--pfm out.pfm] [--aov prefix] [--spp N] [-d cuda] [--gpus N [--nccl]]\n"); return 0; }
        else fprintf(stderr, "ignored argument %s\n", a.c_str());
    }
    if (scene.empty()) { fprintf(stderr, "vision-cli: -s <scene file> is required\n"); return 2; }
    if (device != "cuda") { fprintf(stderr, "vision-cli: only -d cuda exists (no CPU backend)\n"); return 2; }
    try {
        if (gpus > 1) {
            std::vector<float> img; vh::MultiGpuTimes t;
            auto p = vh::render_on_gpus(scene, gpus, spp, nccl, img, &t, dev_index);
            uint32_t n = spp ? spp : p->records.spp;
            fprintf(stderr, "[vision] %ux%u, %u spp over %u GPUs (%s), init %.1f ms, render %.1f ms (%.1f Msamples/s), reduce + resolve + download %.2f ms\n",
                    p->records.width, p->records.height, n, gpus, nccl ? "ncclAllReduce" : "peer-memory reduce",
                    t.prepare_ms, t.render_ms, 1e-3 * p->records.width * p->records.height * n / t.render_ms, t.reduce_ms);
            vh::save_image(out, img.data(), p->records.width, p->records.height);
            if (!pfm.empty()) {
                std::vector<float> acc((size_t)p->records.width * p->records.height * 4);
                p->check(vb_film_download(p->ctx, acc.data(), VB_FILM_ACCUM), "vb_film_download");
                for (size_t i = 0; i < acc.size(); i += 4) {      // SUM film -> mean, like the single-GPU --pfm
                    float inv = acc[i + 3] > 0.f ? 1.f / acc[i + 3] : 0.f;
                    acc[i] *= inv; acc[i + 1] *= inv; acc[i + 2] *= inv; acc[i + 3] = 1.f;
                }
                vh::save_pfm(pfm, acc.data(), p->records.width, p->records.height);
            }
            return 0;
        }
        auto t0 = std::chrono::steady_clock::now();
        auto p = vh::import_scene(scene);
        p->options = options;
        p->prepare(dev_index);
        for (auto &l : p->scene.log) fprintf(stderr, "[vision] %s\n", l.c_str());
        uint32_t n = spp ? spp : p->records.spp;
        auto t1 = std::chrono::steady_clock::now();
        p->render(n);
        std::vector<float> img;
        p->final_picture(img);
        auto t2 = std::chrono::steady_clock::now();
        fprintf(stderr, "[vision] %ux%u, %u spp, %llu triangles, %llu wide nodes, BVH build %.2f ms, init %.1f ms, render %.1f ms\n",
                p->records.width, p->records.height, n, (unsigned long long)p->build_stats.n_triangles,
                (unsigned long long)p->build_stats.n_nodes, p->build_stats.build_ms,
                std::chrono::duration<double, std::milli>(t1 - t0).count(), std::chrono::duration<double, std::milli>(t2 - t1).count());
        vh::save_image(out, img.data(), p->records.width, p->records.height, p->output_is_display_referred());      // .png / .ppm (display gamma) or .hdr / .pfm (linear)
        if (!pfm.empty()) {
            std::vector<float> acc((size_t)p->records.width * p->records.height * 4);
            p->check(vb_film_download(p->ctx, acc.data(), VB_FILM_ACCUM), "vb_film_download");
            vh::save_pfm(pfm, acc.data(), p->records.width, p->records.height);
        }
        if (!aov.empty()) {       // rt_geom's G-buffer planes of frame 0 (the denoiser inputs): <prefix>_normal.pfm, _albedo.pfm, _emission.pfm, _geom.pfm, _motion.pfm, _gradient.pfm
            std::vector<float> plane((size_t)p->records.width * p->records.height * 4);
            const char *names[VB_AOV_COUNT] = {"normal", "albedo", "emission", "geom", "motion", "gradient"};
            p->check(vb_gbuffer_render(p->ctx, 0), "vb_gbuffer_render");
            p->check(vb_gbuffer_gradient(p->ctx), "vb_gbuffer_gradient");
            for (int k = 0; k < VB_AOV_COUNT; ++k) {
                p->check(vb_gbuffer_download(p->ctx, k, plane.data()), "vb_gbuffer_download");
                vh::save_pfm(aov + "_" + names[k] + ".pfm", plane.data(), p->records.width, p->records.height);
            }
        }
    } catch (const std::exception &e) {
        fprintf(stderr, "vision-cli: %s\n", e.what());
        return 1;
    }
    return 0;
}
