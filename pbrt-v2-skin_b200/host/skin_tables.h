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

// Spectrum::FromRGB((1,1,1), SPECTRUM_REFLECTANCE) (core/spectrum.cpp): what "color Kr" [1 1 1] becomes.  Exported
// from the reference build (rgb_unit30 in tests/golden/skin_ld4.npz).  A grey value v converts to (v * white) * .94,
// which equals v * k[i] up to one rounding.
static const float kUnitRGBToSpectrum30[30] = {
    0.997982085f, 0.998490036f, 0.998524427f, 0.99871248f, 0.998672247f,
    0.998708844f, 0.998744309f, 0.998709857f, 0.99849236f, 0.998001933f,
    0.997532845f, 0.997481167f, 0.997701705f, 0.997843564f, 0.998324633f,
    0.998719871f, 0.998711884f, 0.998748541f, 0.998712003f, 0.998708963f,
    0.998764694f, 0.998792112f, 0.998735189f, 0.998644888f, 0.998655975f,
    0.998574495f, 0.99752152f, 0.996186376f, 0.996272445f, 0.99656117f
};

// Molar extinction of oxy-/deoxy-haemoglobin [cm^-1/(mol/L)] at 400..700 nm in 5 nm steps (61 samples;
// Prahl's compilation, as tabulated by the reference in materials/skincoeffs.cpp:35-73).  Data, not code.
static const int kPigmentSamples = 61;
static const float kOxyHaemoglobin61[61] = {
    266200.f, 331450.f, 466800.f, 523100.f, 480400.f, 351100.f, 246100.f, 149050.f,
    102600.f, 78880.f, 62820.f, 51525.f, 44480.f, 38440.f, 33210.f, 29480.f,
    26630.f, 24925.f, 23680.f, 22155.f, 20930.f, 20185.f, 20040.f, 20715.f,
    24200.f, 30885.f, 39960.f, 48335.f, 53240.f, 50985.f, 43020.f, 35650.f,
    32610.f, 35210.f, 44500.f, 54425.f, 50100.f, 30620.f, 14400.f, 6681.5f,
    3200.f, 1958.5f, 1506.f, 1166.5f, 942.f, 740.8f, 610.f, 495.6f,
    442.f, 397.7f, 368.f, 340.3f, 319.6f, 305.6f, 294.f, 283.8f,
    277.6f, 273.6f, 276.f, 280.6f, 290.f,
};
static const float kDeoxyHaemoglobin61[61] = {
    223300.f, 261950.f, 304000.f, 353200.f, 407600.f, 471500.f, 528600.f, 549600.f,
    413300.f, 259950.f, 103300.f, 33435.f, 23390.f, 18700.f, 16160.f, 14920.f,
    14550.f, 15375.f, 16680.f, 18650.f, 20860.f, 23285.f, 25770.f, 28680.f,
    31590.f, 35170.f, 39040.f, 42840.f, 46590.f, 50490.f, 53410.f, 54530.f,
    53790.f, 49700.f, 45070.f, 40905.f, 37020.f, 33590.f, 28320.f, 21185.f,
    14680.f, 12040.f, 9444.f, 7553.5f, 6510.f, 5763.5f, 5149.f, 4666.5f,
    4345.f, 4026.5f, 3750.f, 3481.5f, 3227.f, 3011.f, 2795.f, 2591.f,
    2408.f, 2224.5f, 2052.f, 1923.5f, 1794.f,
};
