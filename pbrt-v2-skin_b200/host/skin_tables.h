// skin_tables.h -- spectral data constants of the reference's colour pipeline (data, not code).
// kGreyRGBToSpectrum30[i] = rgbRefl2SpectWhite averaged over wavelength bin i, times the .94
// of Spectrum::FromRGB (core/spectrum.cpp): a grey RGB value v converts to v * k[i].
// Exported once from the reference build (lights.bin Lemit / 3200, oracle/refbuild/ref_driver.cpp).
#pragma once
static const float kGreyRGBToSpectrum30[30] = {
    0.997982025f, 0.998489976f, 0.998524487f, 0.99871248f, 0.998672247f,
    0.998708904f, 0.998744249f, 0.998709917f, 0.99849236f, 0.998001933f,
    0.997532785f, 0.997481167f, 0.997701705f, 0.997843564f, 0.998324633f,
    0.998719871f, 0.998711944f, 0.998748541f, 0.998712003f, 0.998708963f,
    0.998764753f, 0.998792171f, 0.998735189f, 0.998644888f, 0.998656034f,
    0.998574495f, 0.99752152f, 0.996186376f, 0.996272445f, 0.996561229f
};
