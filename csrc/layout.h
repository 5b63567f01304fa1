// Flat parameter layout of the CIFAR-10 CNN, shared by every kernel.
// Mirrors distributed_neural_network_b200/ops/flat.py (tests/test_flat_and_model.py checks
// that both agree). Tensor order/shapes: reference models/model.py:13-18.
#pragma once

namespace dnnb200 {

constexpr int kAlign = 32;  // floats (128 B)

constexpr int kOffW1 = 0;       // conv1.weight (6,3,5,5)   450
constexpr int kOffB1 = 480;     // conv1.bias   (6)
constexpr int kOffW2 = 512;     // conv2.weight (16,6,5,5)  2400
constexpr int kOffB2 = 2912;    // conv2.bias   (16)
constexpr int kOffF1W = 2944;   // fc1.weight   (120,400)   48000
constexpr int kOffF1B = 50944;  // fc1.bias     (120)
constexpr int kOffF2W = 51072;  // fc2.weight   (84,120)    10080
constexpr int kOffF2B = 61152;  // fc2.bias     (84)
constexpr int kOffF3W = 61248;  // fc3.weight   (10,84)     840
constexpr int kOffF3B = 62112;  // fc3.bias     (10)
constexpr int kFlatNumel = 62144;
constexpr int kParamNumel = 62006;

// The conv part of the gradient is produced per image by the forward/backward
// kernel and summed over the batch by the update kernel; it occupies the prefix
// [0, kConvPrefix) of the flat index space.
constexpr int kConvPrefix = kOffF1W;  // 2944

// Per-image record the forward/backward kernel leaves for the update kernel
// (inputs and deltas of the three FC layers).
constexpr int kSavP2 = 0;     // 400  pooled conv2 output == fc1 input (C,H,W order)
constexpr int kSavH1 = 400;   // 120  relu(fc1)
constexpr int kSavH2 = 520;   // 84   relu(fc2)
constexpr int kSavD1 = 608;   // 120  dLoss/d(fc1 pre-activation)
constexpr int kSavD2 = 728;   // 84   dLoss/d(fc2 pre-activation)
constexpr int kSavD3 = 812;   // 10   dLoss/d(logits)
constexpr int kSavStride = 832;

// Image geometry
constexpr int kInC = 3, kInH = 32, kInW = 32;
constexpr int kImgNumel = kInC * kInH * kInW;  // 3072
constexpr int kC1 = 6, kC1H = 28, kP1H = 14;   // conv1 out 6x28x28, pooled 6x14x14
constexpr int kC2 = 16, kC2H = 10, kP2H = 5;   // conv2 out 16x10x10, pooled 16x5x5
constexpr int kF1 = 120, kF2 = 84, kF3 = 10, kFlat = 400;

// Tensor-core instantiation of the step kernel: the conv weights as the three canonical K-major B operands of its
// implicit GEMMs (fused_cluster.cu), kept in global memory next to the flat parameters and refreshed by the update
// kernel with every step, so that the step kernel fetches them with bulk copies instead of re-packing 5 250 weights
// in shared memory at every launch. Element (row n, K-chunk q, component c) of an operand with NCH chunks sits at
// ((n / 8) * NCH + q) * 32 + (n % 8) * 4 + c; everything that is not a weight is zero and stays zero.
constexpr int kB1Chunks = 26, kB2Chunks = 50, kBdChunks = 100;
constexpr int kPackB1 = 0;                              // conv1.weight: n = o, chunk = tap, c = input channel
constexpr int kPackB2 = kPackB1 + 2 * kB1Chunks * 32;   // conv2.weight: n = o, chunk = 2*tap + ch/4, c = ch%4
constexpr int kPackBd = kPackB2 + 2 * kB2Chunks * 32;   // conv2.weight for the data gradient: n = ch, chunk = 4*(24-tap) + o/4, c = o%4
constexpr int kPackNumel = kPackBd + 2 * kBdChunks * 32;   // 11 264 floats

// Gradient exchange: the flat vector is cut into slices of kSlice floats; one
// CTA of the update kernel owns one slice end to end (wgrad -> push to peers ->
// reduce -> SGD), so the NVLink transfer of early slices overlaps the math of
// later ones.
constexpr int kSlice = 512;
constexpr int kNumSlices = (kFlatNumel + kSlice - 1) / kSlice;  // 122 (last one is short: 62144 = 121*512 + 192)
constexpr int kMaxWorld = 16;

}  // namespace dnnb200
