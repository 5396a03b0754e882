// Standalone mirror of reference include/FEM/EquationParameters.hpp:4-49 (same struct names, fields and defaults)
// without the Eigen dependency: `velocity` is a std::array<float, 3>, indexed with [] like Eigen::Vector3f.
#pragma once
#include <array>

struct EquationParameters {
    float time_step;
};

struct HeatParameters : public EquationParameters {  // reference :8-15
    float conductivity = 0.05f;
    HeatParameters() { time_step = 0.01f; }
};

struct AdvectionDiffusionParameters : public EquationParameters {  // reference :17-27
    float c = 0.25f;
    std::array<float, 3> velocity{{1.0f, 0.0f, 0.0f}};
    AdvectionDiffusionParameters() { time_step = 0.001f; }
};

struct WaveParameters : public EquationParameters {  // reference :29-36
    float c = 0.05f;
    WaveParameters() { time_step = 0.05f; }
};

struct ReactionDiffusionParameters : public EquationParameters {  // reference :38-49
    float Du = 0.08f;
    float Dv = 0.04f;
    float feed_rate = 0.035f;
    float kill_rate = 0.06f;
    ReactionDiffusionParameters() { time_step = 0.001f; }
};
