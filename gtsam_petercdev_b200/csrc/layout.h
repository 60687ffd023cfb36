// layout.h -- sizes of the rows the C ABI exchanges (include/gtsam_b200.h): state rows per variable type, measurement rows,
// residual dimensions and key counts per factor bucket.  Shared by the engine and by the layers written on top of its ABI.
#pragma once

namespace gtb {

static const int kZIn[7] = {3, 12, 2, 3, 12, 17, 3};       // measurement doubles per factor as given by the caller
static const int kZDev[7] = {4, 12, 2, 4, 12, 17, 3};      // ... as kept on the device (Pose2 -> cos, sin, x, y)
static const int kM[7] = {3, 6, 2, 3, 6, 9, 3};            // residual dimension
static const int kW[7] = {7, 13, 13, 4, 7, 10, 4};         // columns of the whitened [A|b] row block
static const int kNKeys[7] = {2, 2, 2, 1, 1, 1, 1};
static const int kVT[7][2] = {{0, 0}, {1, 1}, {2, 3}, {0, -1}, {1, -1}, {2, -1}, {3, -1}};   // variable types of the keys
static const int kStateIn[4] = {3, 12, 17, 3};             // state doubles per variable as given by the caller
static const int kStateDev[4] = {4, 12, 17, 3};            // ... as kept on the device and returned by gtb_get_variables
static const int kDim[4] = {3, 6, 9, 3};                   // tangent dimension

}  // namespace gtb
