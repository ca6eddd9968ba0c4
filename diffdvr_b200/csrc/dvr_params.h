// Kernel-argument PODs shared by the host dispatch (dvr_capi.cpp) and the sm_100a kernels.
// They replace the reference's by-value accessor structs (kernel::RendererInputs & co,
// renderer/renderer_settings.cuh:150-244): strides are pre-resolved on the host (size-1
// dimensions -> stride 0, renderer/renderer_tensor.cuh:140-144), scalar variants travel by
// value, and the reference-frame tangents are computed once on the host.
#pragma once
#include <stdint.h>

// Layout of the QUAD buffers (RayInputs::quads, AdjointArgs::gradQuads): 0 = Q[x][y][z] (z innermost, the two quads of
// a cell Y*Z entries apart), 1 = Q[y][z][x] (x innermost: the two quads of a cell are NEIGHBOURS -- one address, an
// immediate +16 for the second access, and 7 of 8 cells have both quads in one 128-byte line)
#ifndef DVR_QUAD_XINNER
#define DVR_QUAD_XINNER 1
#endif

namespace dvr {

// tile geometry: one warp = 8x4 pixels, one CTA = 2x2 warps = 16x8 pixels
constexpr int kWarpW = 8, kWarpH = 4;
constexpr int kCtaW = 16, kCtaH = 8;
constexpr int kThreads = 128;

// numeric enum values of renderer/renderer_settings.cuh:81-146
enum : int { kFilterNearest = 0, kFilterTrilinear = 1, kFilterPreshaded = 3 };
// internal: trilinear sampling of the library's own QUAD copy of the volume (RayInputs::quads), chosen by the
// host dispatch (dvr_capi.cu); never part of the C-ABI
enum : int { kFilterTrilinearQuad = 101 };
enum : int { kCamRays = 0, kCamInvView = 1, kCamFrame = 2 };
enum : int { kTfIdentity = 0, kTfTexture = 1, kTfLinear = 2, kTfGaussian = 3, kTfPreshaded = 5, kTfGaussianLog = 6 };
enum : int { kBlendBeer = 0, kBlendAlpha = 1 };

struct Strided1 { const float* p; long long b, c; };             // (B)*3
struct Strided3 { const float* p; long long b, y, x; };          // (B)*(H)*(W)
struct Strided4 { const float* p; long long b, y, x, c; };       // (B)*H*W*3|4
struct StridedMat { const float* p; long long b, r, c; };        // (B)*R*C

struct RayInputs {
  int W, H, B;
  int X, Y, Z;                       // volume resolution
  // volume, element offsets: base + b*volB + x*sx + y*sy + z*sz (per-batch span < 2^31)
  const float* volume;
  long long volB;
  int sx, sy, sz;
  // QUAD copy (kFilterTrilinearQuad): quad (x, y, z) = (v[y][z], v[y][z'], v[y'][z], v[y'][z']) of slice min(x, X-1)
  // for x in [0, X] (X + 1 slices: the far-face clamp of x is a duplicated last slice), y' = min(y+1, Y-1),
  // z' = min(z+1, Z-1), stored as Q[b][y][z][x] (DVR_QUAD_XINNER, the default: one cell = 2 aligned 16-byte loads at
  // quad offsets q and q + 1) or Q[b][x][y][z] (DVR_QUAD_XINNER = 0: q and q + Y*Z); no clamp in the kernel
  const float4* quads;
  long long quadB;                   // float4 elements per batch entry, (X+1)*Y*Z (0: one volume for all views)
  // hoisted index arithmetic of the QUAD cell load: q = ux*quadYZ + uy*Z + uz - quadBias on the biased-float bit
  // patterns of the cell coordinates (dvr_device.cuh load_cell), clamped to quadMax = X*Y*Z - 1 as a safety net
  unsigned quadYZ, quadBias, quadMax;
  // DVR_QUAD_XINNER layout Q[b][y][z][x] (x in [0, X], innermost): q = uy*quadYZ + uz*quadSZ + ux - quadBias with
  // quadYZ = Z*(X+1) (the y stride), quadSZ = X+1, quadMax = (X+1)*Y*Z - 2; the x + 1 quad of a cell is q + 1
  unsigned quadSZ;
  Strided1 boxMin, boxSize;          // p == nullptr -> use the *V value
  float boxMinV[3], boxSizeV[3];
  int cameraMode;
  Strided4 rayStart, rayDir;         // kCamRays
  StridedMat camMat;                 // kCamInvView (4x4) / kCamFrame (3x3); p == nullptr -> camMatV
  float camMatV[16];
  float tanFovX, tanFovY;            // kCamFrame, computed on the host (tanf of the half angles)
  Strided3 stepSize;                 // p == nullptr -> stepV
  float stepV;
  StridedMat tf;                     // (B)*R*C
  int R, C;
  float earlyOut;
};

struct ForwardOutputs {
  float* color;                      // B*H*W*4
  long long cB, cY, cX, cC;
  int colorVec4;                     // 1: cC==1, 16-byte aligned rows -> one STG.128 per pixel
  int* term;                         // B*H*W
  long long tB, tY, tX;
};

struct AdjointArgs {
  // outputs of the forward pass + adjoint colour
  Strided4 color;                    // (B)*H*W*4
  const int* term;
  long long tB, tY, tX;
  Strided4 adjColor;
  // outputs (accumulated)
  float* adjVolume;                  // element offsets like the volume
  long long avB;
  int avx, avy, avz;
  int sameStrides;                   // adj_volume has the element strides of the volume
  // kFilterTrilinearQuad + hasVolume: gradients are accumulated into a zeroed QUAD scratch
  // G[b][x][y][z] = (g(y,z), g(y,z'), g(y',z), g(y',z')) with 2 x red.v4 per cell and folded into adjVolume
  // by k_fold_quads afterwards
  float4* gradQuads;                 // (X+1)*Y*Z quads per batch entry, like RayInputs::quads
  long long gradQuadB;               // float4 elements per batch entry (0: one gradient volume for all views)
  float* adjStep;
  long long asB, asY, asX;
  int stepAtomic;                    // broadcasting -> atomics (renderer_kernels.cuh:418-421)
  float* adjRayStart;
  long long rsB, rsY, rsX, rsC;
  float* adjRayDir;
  long long rdB, rdY, rdX, rdC;
  int camAtomic;
  float* adjTf;
  long long atB, atR, atC;
  int hasStep, hasCam;
  int compatStepOverwrite;           // see include/dvr_b200.h
  int marchZeroRays;                 // 1: march rays whose adjoint colour is all zero too (DvrCallOptions)
};

struct ForwardGradArgs {
  int D;                             // total number of tangents
  int d0;                            // first tangent slot handled by this pass
  int writeColor;                    // only the first pass writes colour / terminationIndex
  int dStep;
  int dRayStart[3], dRayDir[3];
  const int* dTf;                    // (B)*R*C slots or nullptr
  long long dtB, dtR, dtC;
  float* grad;                       // B*H*W*D*4
  long long gB, gY, gX, gD, gC;
  int exactPos;                      // k_tf_jacobian: sample positions in the reference's operation order (when the
                                     // colour it reads was written by k_forward_gradients_inputs)
  int gradVec4;                      // 1: gC == 1, every other stride a multiple of 4, 16-byte aligned base -> the four
                                     // channels of a slot are one 16-byte access
};

struct FwdVarsArgs {
  int W, H, B, D;
  const float* fv;                   // B*H*W*D*4
  long long fB, fY, fX, fD, fC;
  Strided4 gradColor;
  int dStep;
  int dRayStart[3], dRayDir[3];
  const int* dTf;
  long long dtB, dtR, dtC;
  int R, C, hasTf;
  float* adjStep;
  long long asB, asY, asX;
  float* adjRayStart;
  long long rsB, rsY, rsX, rsC;
  float* adjRayDir;
  long long rdB, rdY, rdX, rdC;
  float* adjTf;
  long long atB, atR, atC;
};

// Camera::generateRays forward / adjoint (renderer/camera_impl.cu:17-94, 96-356)
struct GenRaysArgs {
  int W, H, B;
  StridedMat viewport;               // B*3*3 (eye, right, up)
  float tanFovX, tanFovY;
  float* rayStart;                   // B*H*W*3 (forward: written; adjoint: read)
  long long sB, sY, sX, sC;
  float* rayDir;
  long long dB, dY, dX, dC;
  int packed;                        // both ray tensors contiguous, W % 4 == 0, 16-byte aligned
  float* adjViewport;                // adjoint: B*3*3, zeroed by the launcher, accumulated
  long long avB, avR, avC;
};

// Renderer::compareToImage (renderer/renderer_compareToImage.cuh:9-65, renderer_cuda_impl.cu:9-172)
struct CompareArgs {
  int W, H, B, D;
  Strided4 color, reference;         // B*H*W*4
  const float* grad;                 // B*H*W*D*4
  long long gB, gY, gX, gD, gC;
  int reduce;
  float normalization;               // 1 / (B*H*W), reduce only
  float* diffOut;                    // !reduce: B*H*W*1 ; reduce: 1
  long long oB, oY, oX;
  float* gradOut;                    // !reduce: B*H*W*D ; reduce: D
  long long qB, qY, qX, qD;
};

// TFUtils::preshadeVolume forward / adjoint (renderer/tf_utils_impl.cu:16-92, 98-183). The density volume and
// the TF travel in a RayInputs (volume, X/Y/Z, strides, tf, R, C; everything else unused).
struct PreshadeArgs {
  float* color;                      // forward: 4*X*Y*Z written; adjoint: adj_color read
  long long cC, cX, cY, cZ;
  float* adjVolume;                  // adjoint: 1*X*Y*Z, overwritten
  long long ax, ay, az;
};

// Volume::ImplicitEquation (renderer/volume.h:163-181)
enum : int {
  kEqMarschnerLobb = 0, kEqCube, kEqSphere, kEqInverseSphere, kEqDingDong, kEqEndrass, kEqBarth, kEqHeart, kEqKleine,
  kEqCassini, kEqSteiner, kEqCrossCap, kEqKummer, kEqBlobby, kEqTube, kEqCount
};

// Volume::createImplicitDataset (renderer/volume_implicit.cpp:224-256 -> volume.cpp:369-387)
struct ImplicitArgs {
  int equation;
  int X, Y, Z;
  float boxMin, scaleStep;           // position = boxMin + index * scaleStep
  float fM, alpha, scale, a;         // MarschnerLobb "fM", "alpha"; Cube "scale"; Cassini "a"
  float* out;
  long long sx, sy, sz;
};

// .cvol body (X fastest; u8 | u16 | f32) <-> the fp32 density tensor (renderer/volume.cpp:86-151, 232-276)
struct RawVolumeArgs {
  void* raw;
  int X, Y, Z;
  float* vol;
  long long sx, sy, sz;
};

// Volume::createMipmapLevel (renderer/volume.cpp:359-365)
struct DownsampleArgs {
  const float* in;
  int X, Y, Z;
  long long sx, sy, sz;
  float* out;
  int oX, oY, oZ;
  long long ox, oy, oz;
};

}  // namespace dvr
