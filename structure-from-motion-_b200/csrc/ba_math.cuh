// ba_math.cuh -- per-observation arithmetic of the bundle-adjustment chain (filled in with K8-K12).
#pragma once
#include "lm_small.cuh"
#include "pose_math.cuh"
