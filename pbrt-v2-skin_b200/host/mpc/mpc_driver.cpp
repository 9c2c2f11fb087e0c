// mpc_driver.cpp -- calls the MPC_* API the way the reference's call sites do (core/multipole.cpp:241-295 MultipoleProfileTask::Run:
// fill MPC_LayerSpec[] and MPC_Options, MPC_ComputeDiffusionProfile, MPC_ResampleForUniformDistanceSquaredDistribution, read the
// arrays, MPC_FreeOutput; renderers/profilefit.cpp:190-196 adds MPC_ResampleDistribution), through include/mpc_cuda.h only.
//   mpc_driver in.txt out.bin      same line / record format as `ref_driver mpc` (oracle/refbuild/ref_driver.cpp runMPC)
#include <cmath>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>
#include "../../../include/mpc_cuda.h"

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: mpc_driver <in.txt> <out.bin>\n"); return 2; }
    std::ifstream in(argv[1]);
    FILE *out = fopen(argv[2], "wb");
    if (!in || !out) { fprintf(stderr, "mpc_driver: cannot open files\n"); return 2; }
    std::string line;
    int n = 0;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        int nLayers, desiredLength, lerp, resample; float stepSize;
        if (!(ls >> nLayers >> desiredLength >> lerp >> resample >> stepSize)) continue;
        std::vector<MPC_LayerSpec> specs(nLayers);
        float mfpTotal = 0.f;
        for (int l = 0; l < nLayers; ++l) {
            ls >> specs[l].ior >> specs[l].thickness >> specs[l].mua >> specs[l].musp;
            specs[l].g_HG = 0.f;
            mfpTotal += 1.f / (specs[l].mua + specs[l].musp);
        }
        MPC_Options opt; opt.desiredLength = desiredLength; opt.lerpOnThinSlab = lerp != 0;
        float mfp = mfpTotal / (float)nLayers;
        opt.desiredStepSize = stepSize > 0.f ? stepSize : 12.f * mfp / (float)opt.desiredLength;
        MPC_Output *o = NULL;
        MPC_ComputeDiffusionProfile(nLayers, specs.data(), &opt, &o);
        if (!o) { fprintf(stderr, "mpc_driver: MPC_ComputeDiffusionProfile failed: %s\n", MPC_LastError()); return 1; }
        if (resample == 1) MPC_ResampleForUniformDistanceSquaredDistribution(o);
        else if (resample == 3) MPC_ResampleForUniformDistanceSquaredDistribution(o, 1000);
        else if (resample == 2) {
            std::vector<float> r(777);
            float rmax = 1.05f * sqrtf(o->pDistanceSquared[o->length - 1]);
            for (int i = 0; i < 777; ++i) r[i] = (float)i * rmax / 776.f;
            MPC_ResampleDistribution(o, 777, r.data());
        }
        uint32 len = o->length;
        fwrite(&len, 4, 1, out);
        fwrite(&o->totalReflectance, 4, 1, out); fwrite(&o->totalTransmittance, 4, 1, out);
        fwrite(o->pDistanceSquared, 4, len, out); fwrite(o->pReflectance, 4, len, out); fwrite(o->pTransmittance, 4, len, out);
        MPC_FreeOutput(o);
        ++n;
    }
    // MPC_CreateOutput / MPC_FreeOutput pair (mcprofile.cpp:402-423 fills such an output itself)
    MPC_Output *blank = NULL;
    MPC_CreateOutput(16, &blank);
    if (!blank || blank->length != 16) return 1;
    MPC_FreeOutput(blank);
    MPC_ClearCache();
    fclose(out);
    printf("mpc_profiles %d\n", n);
    return 0;
}
