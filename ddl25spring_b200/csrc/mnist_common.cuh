// Shared definitions of the packed MnistCnn training engine (csrc/mnist_*.cu).
//
// G simulated clients train in lock-step on one GPU: every kernel processes the current
// minibatch of ALL clients at once ("grouped"), each client with its own weights taken from
// its row of the [G, P] flat weight arena (FlatArena layout of models/mnist_cnn.py).
//
// Activation layouts (row = g * B + b, B = rows reserved per client):
//   a1     [rows][26*26][32]  fp32 (tf32-rounded), NHWC   conv1 + ReLU output
//   pooled [rows][64][12][12] fp32, CHW (= torch flatten order, so fc1.weight needs no permute)
//   pmask  [rows][9216] u8: bits 0-1 argmax inside the 2x2 window, bit 2 = gradient flows
//   h1     [rows][128]        fc1 + ReLU output (before dropout2)
//   dz1    [rows][128]        grad wrt fc1 pre-activation
//   dpool  [rows][9216]       grad wrt pooled
//   dy2    [rows][26*26][64]  grad wrt conv2 pre-activation, NHWC on the 26-pitch grid, zero where
//                             h >= 24 or w >= 24 (the flat "shifted window" trick needs those zeros)
//   da1    [rows][26*26][32]  grad wrt conv1 pre-activation (ReLU mask applied)
#pragma once
#include <cstdint>

namespace mnist {

constexpr int P = 1200000;          // padded arena length (1 199 882 parameters)
constexpr int OFF_C1W = 0;          // conv1.weight [32][1][3][3]
constexpr int OFF_C1B = 288;        // conv1.bias   [32]
constexpr int OFF_C2W = 320;        // conv2.weight [64][32][3][3]
constexpr int OFF_C2B = 18752;      // conv2.bias   [64]
constexpr int OFF_F1W = 18816;      // fc1.weight   [128][9216]
constexpr int OFF_F1B = 1198464;    // fc1.bias     [128]
constexpr int OFF_F2W = 1198592;    // fc2.weight   [10][128]
constexpr int OFF_F2B = 1199872;    // fc2.bias     [10]

constexpr int IMG = 28, H1 = 26, H2 = 24, HP = 12;
constexpr int PIX1 = H1 * H1;       // 676 conv1 output positions == flat grid of conv2
constexpr int C1 = 32, C2 = 64;
constexpr int FLAT = C2 * HP * HP;  // 9216
constexpr int FC1 = 128, NCLS = 10;
constexpr int MAX_G = 16;           // clients per launch

// conv2 implicit GEMM on the flat 26-pitch grid: output tile = 128 consecutive flat positions
// starting at row h0 (multiple of 4) of one sample; tap (r, s) reads input rows shifted by
// r*26 + s, all served from one staged window of WIN_ROWS rows.
constexpr int TILE_M = 128;
constexpr int WIN_ROWS = 184;       // 128 + 2*26 + 2 = 182, rounded up to a multiple of 8
constexpr int TILES_PER_SAMPLE = 6; // 24 output rows / 4

struct StepInfo {
  int G;                 // clients in this launch
  int B;                 // rows reserved per client in the activation buffers
  int bsz[MAX_G];        // valid rows of each client in this step (0 = client idle)
  uint32_t seed;         // dropout stream of this step
  float p1, p2;          // dropout probabilities (0 in eval mode)
  float lr;
};

}  // namespace mnist
