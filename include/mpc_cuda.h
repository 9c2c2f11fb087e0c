/* mpc_cuda.h -- the reference's MPC_* profile-calculator API, served by the B200 device layer.
 *
 * Drop-in for src/multipole/MultipoleProfileCalculator/MultipoleProfileCalculator.h:34-76 (PODs :34-67, functions :70-76):
 * same names, same C++ linkage (the reference's header is not extern "C" and MPC_ResampleFor... has a default argument),
 * same ownership -- the callee allocates MPC_Output and its arrays with new / new[], the caller releases them with
 * MPC_FreeOutput; the resample functions replace the arrays in place.  The three call sites that bind the DLL today
 * (core/multipole.cpp:268-291, renderers/mcprofile.cpp:402-423, renderers/profilefit.cpp:190-196) compile against this header
 * unchanged and link libMultipoleProfileCalculator.so (pbrt-v2-skin_b200/Makefile) instead of the reference's DLL.
 *
 * Behind it: skc_profile_batch / skc_profile_resample of include/skin_cuda.h on a lazily created, process-wide device context
 * (the current CUDA device at the first call).  Calls from several threads are serialised on that context -- the reference's
 * 30 MultipoleProfileTasks may call concurrently.  Errors: the reference's functions return void and fail silently
 * (`if (!oppOutput || !numLayers) return;`, MultipoleProfileCalculator.cpp:297); here a device failure leaves *oppOutput NULL
 * (compute) or the output untouched (resample) and MPC_LastError() says why.  There is no CPU fallback.
 */
#ifndef MPC_CUDA_H
#define MPC_CUDA_H

typedef unsigned int uint32;      /* mpc-types.h:40 */

typedef struct {
    float ior;          /* index of refraction */
    float thickness;    /* layer thickness */
    float mua;          /* absorption coefficient */
    float musp;         /* reduced scattering coefficient */
    float g_HG;         /* anisotropy (unused by the calculator) */
} MPC_LayerSpec;

typedef struct {
    float desiredStepSize;   /* distance between adjacent samples */
    uint32 desiredLength;    /* samples in one direction */
    bool lerpOnThinSlab;     /* lerp when the slab is thinner than 2 mfp */
} MPC_Options;

typedef struct {
    uint32 length;
    float *pDistanceSquared;
    float *pReflectance;
    float *pTransmittance;
    float totalReflectance;
    float totalTransmittance;
} MPC_Output;

void MPC_CreateOutput(uint32 length, MPC_Output **oppOutput);
void MPC_ComputeDiffusionProfile(uint32 numLayers, const MPC_LayerSpec *pLayerSpecs, const MPC_Options *pOptions, MPC_Output **oppOutput);
void MPC_ResampleForUniformDistanceSquaredDistribution(MPC_Output *pOutput, uint32 targetLength = 0);
void MPC_ResampleDistribution(MPC_Output *pOutput, uint32 nSamplePoints, float *samplePoints);
void MPC_FreeOutput(MPC_Output *pOutput);
void MPC_ClearCache();     /* releases the device context and everything it caches */

/* Not in the reference: the message of the last failed call of this thread's process ("" if none). */
const char *MPC_LastError();

#endif /* MPC_CUDA_H */
