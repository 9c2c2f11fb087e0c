// skin_material.h -- host side of the LayeredSkin material: pigment model -> per-wavelength layer
// coefficients and the batched device call that turns them into the Rd
// profile tables the Mo kernel samples (row a10's caller, core/multipole.cpp:241-295,396-420).
#pragma once
#include <cstdint>
#include <vector>
#include "skin_scene.h"

namespace skin {

static const int kWavelengths = 30;       // nSpectralSamples (core/spectrum.h:44-46), 400..700 nm

// what LayeredSkin's constructor hands to ComputeMultipoleProfile (materials/layeredskin.cpp:49-92)
struct SkinLayerSpectra {
    float mua[2][kWavelengths];           // absorption, scene units^-1
    float musp[2][kWavelengths];          // reduced scattering
    float eta[2], thickness[2];
};

// materials/skincoeffs.h:46-135 + core/material.h:95-156 + SampledSpectrum::FromSampled
SkinLayerSpectra ComputeSkinLayerSpectra(const SkinParams &p);

// Profile (core/multipole.h) for all wavelengths, table-major: rd[w*len + k]
struct ProfileTable {
    uint32_t len = 0;
    std::vector<float> rd;
    float dsqSpacing[kWavelengths], rcpDsqSpacing[kWavelengths];
    float totalReflectance[kWavelengths], totalTransmittance[kWavelengths];
};

// MPC_LayerSpec / MPC_Options for every wavelength (MultipoleProfileTask::Run, core/multipole.cpp:241-268)
void BuildProfileJobs(const SkinLayerSpectra &s, int desiredLength, bool lerpOnThinSlab,
                      std::vector<skc_layer_spec> &specs, std::vector<uint32_t> &layersPerSet,
                      std::vector<skc_mpc_options> &opts);

// ComputeMultipoleProfile (core/multipole.cpp:396-420): one skc_profile_batch call for the 30 wavelengths -- or, with a
// communicator, for this rank's share of them (wavelength w on rank w % world), the finished tables all-gathered.
// Returns 0 or the device error code (skc_last_error(ctx) has the text).
int ComputeMultipoleProfile(skc_ctx *ctx, const SkinLayerSpectra &s, int desiredLength, bool lerpOnThinSlab,
                            ProfileTable &out, skc_comm *comm = nullptr);

}  // namespace skin
