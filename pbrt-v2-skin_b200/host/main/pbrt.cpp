// pbrt.cpp -- the `pbrt` executable of the device path: same command line as the reference's main/pbrt.cpp:68-115
//   pbrt [--ncores n] [--outfile filename] [--quick] [--quiet] [--verbose] [--help] <filename.pbrt> ...
// plus --device n (CUDA device ordinal, default 0) and --gpus N (render each frame on N GPUs, devices n .. n + N - 1: irradiance
// points and camera tiles sharded, NCCL all-gather / film reduce).  --ncores does not set a thread count here (the work runs on the
// GPU); it keeps its second meaning in the reference: the task counts that seed the per-task random streams
// (multipolesubsurface.cpp:194-199, samplerrenderer.cpp:206-208), so a frame rendered with the same --ncores draws the
// same samples as the reference does.
#include "../skin_scene.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

int main(int argc, char *argv[]) {
    skin::LoadOptions opt;
    int nCores = 0, device = 0, nGpus = 1;
    bool quiet = false, verbose = false;
    std::vector<std::string> filenames;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--ncores") && i + 1 < argc) nCores = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--outfile") && i + 1 < argc) opt.imageFile = argv[++i];
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--gpus") && i + 1 < argc) nGpus = std::max(1, atoi(argv[++i]));
        else if (!strcmp(argv[i], "--quick")) opt.quickRender = true;
        else if (!strcmp(argv[i], "--quiet")) quiet = true;
        else if (!strcmp(argv[i], "--verbose")) verbose = true;
        else if (!strcmp(argv[i], "--foreground")) {}
        else if (!strcmp(argv[i], "--help") || !strcmp(argv[i], "-h")) {
            printf("usage: pbrt [--ncores n] [--outfile filename] [--quick] [--quiet] [--verbose] [--help] [--device n] [--gpus N] <filename.pbrt> ...\n");
            return 0;
        } else filenames.push_back(argv[i]);
    }
    if (nCores <= 0) nCores = (int)std::max(1u, std::thread::hardware_concurrency());      // NumSystemCores (core/parallel.cpp)
    if (!quiet) printf("pbrt (B200 subsurface path) [task streams seeded for %d core(s)]\n", nCores);
    if (filenames.empty()) { fprintf(stderr, "pbrt: no scene file given (scene descriptions on standard input are not supported)\n"); return 1; }
    skc_ctx *ctx = skc_create(device);
    if (!ctx) { fprintf(stderr, "pbrt: %s\n", skc_last_error(nullptr)); return 1; }
    int rc = 0;
    for (const std::string &f : filenames) {
        skin::SceneDesc sd;
        if (!sd.Load(f, &opt)) { fprintf(stderr, "Error: %s\n", sd.error.c_str()); rc = 1; continue; }
        if (!quiet) for (const std::string &w : sd.warnings) fprintf(stderr, "Warning: %s\n", w.c_str());
        std::vector<float> rgb;
        skin::RenderStats st = {};
        std::string err;
        if (sd.rendererName == "tessellatesurfacepoints" || sd.rendererName == "surfacepoints") {
            uint64_t n = 0;
            if (!skin::TessellateSurfacePointsToFile(ctx, sd, &n, &err)) { fprintf(stderr, "Error: %s: %s\n", f.c_str(), err.c_str()); rc = 1; continue; }
            if (!quiet) printf("Obtained %llu surface points%s%s\n", (unsigned long long)n, sd.rendererFilename.empty() ? ". (Not writing to file)" : ", writing to file ",
                               sd.rendererFilename.c_str());
            continue;
        }
        if (!(nGpus > 1 ? skin::RenderSceneMultiGPU(device, nGpus, sd, nCores, &rgb, &st, &err) : skin::RenderScene(ctx, sd, nCores, &rgb, &st, &err))) { fprintf(stderr, "Error: %s: %s\n", f.c_str(), err.c_str()); rc = 1; continue; }
        std::string name = sd.filmFilename.empty() ? "pbrt.tga" : sd.filmFilename;
        if (!skin::WriteImage(name, rgb.data(), sd.xres, sd.yres, &err)) { fprintf(stderr, "Error: %s\n", err.c_str()); rc = 1; continue; }
        if (!quiet)
            printf("%s: %dx%d, %llu irradiance points (%u octree nodes), %llu camera samples; scene %.3f s, material %.3f s, "
                   "Preprocess %.3f s, Rendering %.3f s -> %s\n", f.c_str(), sd.xres, sd.yres, (unsigned long long)st.nPoints, st.octreeNodes,
                   (unsigned long long)st.nSamples, st.buildSec, st.materialSec, st.preprocessSec, st.renderSec, name.c_str());
        (void)verbose;
    }
    skc_destroy(ctx);
    return rc;
}
