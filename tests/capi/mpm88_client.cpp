// tests/capi/mpm88_client.cpp -- the flow of the reference's own C-API test (c_api/tests/mpm88_test.cpp:18-88: ti::Runtime,
// load_aot_module, allocate_ndarray, kernel argument binding, 50 x {reset_grid, p2g, update_grid_v, g2p}, wait) without
// gtest, compiled against the REFERENCE's headers (<taichi/cpp/taichi.hpp>, taichi_core.h, taichi_cuda.h; include/b200mpm.h
// only adds its Part 2 after them) and linked with libb200mpm.so.  Prints one line of results for the Python test.
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>

#include "taichi/cpp/taichi.hpp"
#include "b200mpm.h"  // Part 1 is skipped (TAICHI_H is defined): the b2_* declarations compile against the reference's types

namespace {
constexpr int kNrParticles = 8192 * 2;
constexpr int kNGrid = 128;
constexpr size_t N_ITER = 50;

struct MPM88Demo {
  ti::Runtime runtime_;
  ti::AotModule module_;
  ti::NdArray<float> x_, v_, J_, C_, grid_v_, grid_m_, pos_;
  ti::Kernel k_init_particles_, k_substep_reset_grid_, k_substep_p2g_, k_substep_update_grid_v_, k_substep_g2p_;

  MPM88Demo(const std::string &aot_path, TiArch arch) {
    runtime_ = ti::Runtime(arch);
    module_ = runtime_.load_aot_module(aot_path.c_str());
    const std::vector<uint32_t> shape_1d = {kNrParticles}, shape_2d = {kNGrid, kNGrid};
    const std::vector<uint32_t> vec2_shape = {2}, vec3_shape = {3}, mat2_shape = {2, 2};
    x_ = runtime_.allocate_ndarray<float>(shape_1d, vec2_shape, true);
    v_ = runtime_.allocate_ndarray<float>(shape_1d, vec2_shape);
    pos_ = runtime_.allocate_ndarray<float>(shape_1d, vec3_shape);
    C_ = runtime_.allocate_ndarray<float>(shape_1d, mat2_shape);
    J_ = runtime_.allocate_ndarray<float>(shape_1d, {});
    grid_v_ = runtime_.allocate_ndarray<float>(shape_2d, vec2_shape);
    grid_m_ = runtime_.allocate_ndarray<float>(shape_2d, {});
    k_init_particles_ = module_.get_kernel("init_particles");
    k_substep_g2p_ = module_.get_kernel("substep_g2p");
    k_substep_reset_grid_ = module_.get_kernel("substep_reset_grid");
    k_substep_p2g_ = module_.get_kernel("substep_p2g");
    k_substep_update_grid_v_ = module_.get_kernel("substep_update_grid_v");
    k_init_particles_[0] = x_, k_init_particles_[1] = v_, k_init_particles_[2] = J_;
    k_substep_reset_grid_[0] = grid_v_, k_substep_reset_grid_[1] = grid_m_;
    k_substep_p2g_[0] = x_, k_substep_p2g_[1] = v_, k_substep_p2g_[2] = C_, k_substep_p2g_[3] = J_;
    k_substep_p2g_[4] = grid_v_, k_substep_p2g_[5] = grid_m_;
    k_substep_update_grid_v_[0] = grid_v_, k_substep_update_grid_v_[1] = grid_m_;
    k_substep_g2p_[0] = x_, k_substep_g2p_[1] = v_, k_substep_g2p_[2] = C_, k_substep_g2p_[3] = J_;
    k_substep_g2p_[4] = grid_v_, k_substep_g2p_[5] = pos_;
    k_init_particles_.launch();
    runtime_.wait();
  }
  void Step() {
    for (size_t i = 0; i < N_ITER; i++) {
      k_substep_reset_grid_.launch();
      k_substep_p2g_.launch();
      k_substep_update_grid_v_.launch();
      k_substep_g2p_.launch();
    }
    runtime_.wait();
  }
};
}  // namespace

int main() {
  uint32_t narch = 0;
  ti_get_available_archs(&narch, nullptr);
  std::vector<TiArch> archs(narch);
  ti_get_available_archs(&narch, archs.data());
  bool cuda = false;
  for (TiArch a : archs) cuda = cuda || a == TI_ARCH_CUDA;
  if (!cuda) {
    printf("no cuda arch\n");
    return 2;
  }
  MPM88Demo demo("builtin:mpm88", TI_ARCH_CUDA);
  std::vector<float> x0(kNrParticles * 2), x1(kNrParticles * 2);
  demo.x_.read(x0);
  demo.Step();
  demo.x_.read(x1);
  double my0 = 0, my1 = 0;
  bool finite = true;
  for (int p = 0; p < kNrParticles; p++) {
    my0 += x0[2 * p + 1], my1 += x1[2 * p + 1];
    finite = finite && std::isfinite(x1[2 * p]) && std::isfinite(x1[2 * p + 1]);
  }
  uint32_t ncap = 99;
  ti_get_runtime_capabilities(demo.runtime_, &ncap, nullptr);          // taichi_core.h:951: LLVM archs report none
  const TiAotModule none = ti_create_aot_module(demo.runtime_, "x", 1);  // :1084: TCM bytes are not loadable here
  const TiError e = ti_get_last_error(nullptr, nullptr);
  ti_set_last_error(TI_ERROR_SUCCESS, nullptr);
  // Part 2 through the reference's handle types: the launch counter of the runtime the C++ wrapper owns
  const unsigned long long launches = b2_launch_count(demo.runtime_);
  printf("mpm88_client finite=%d mean_y0=%.6f mean_y1=%.6f capabilities=%u tcm_module=%d tcm_error=%d launches=%llu\n", (int)finite,
         my0 / kNrParticles, my1 / kNrParticles, ncap, none == TI_NULL_HANDLE ? 0 : 1, (int)e, launches);
  return finite && my1 < my0 && ncap == 0 && none == TI_NULL_HANDLE && e == TI_ERROR_NOT_SUPPORTED && launches >= 4 * N_ITER ? 0 : 1;
}
