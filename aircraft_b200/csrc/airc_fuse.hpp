// airc_fuse.hpp -- host-side (float64) folding of the checkpoint and aircraft constants into
// the FP32 constant block the kernels take by value.
//
// Folded from /root/reference/src/aircraft/surrogates/models.py:114-123,143-155:
//   y = (W3 tanh(W2 (W1 z + b1) + b2) + b3) * sigma_out + mu_out,   z = (in - mu_in)/sigma_in
// There is no activation between the first two Linear layers, so W21 = W2 W1 exactly.
#pragma once
#include <math.h>
#include <string.h>

#include "../../include/aircraft_b200.h"
#include "airc_dynamics.cuh"

namespace airc {

inline void fuse_network(const AircNNBlob& nn, FusedNet& out) {
  const double kappa = 2.0 / log(2.0);  // 2*log2(e)
  double W21[32][5], b21[32];
  for (int j = 0; j < 32; ++j) {
    double b = nn.b2[j];
    for (int m = 0; m < 16; ++m) b += (double)nn.W2[j * 16 + m] * (double)nn.b1[m];
    for (int i = 0; i < 5; ++i) {
      double s = 0.0;
      for (int m = 0; m < 16; ++m) s += (double)nn.W2[j * 16 + m] * (double)nn.W1[m * 5 + i];
      W21[j][i] = s;
    }
    b21[j] = b;
  }
  for (int j = 0; j < 32; ++j) {
    double a0 = b21[j];
    for (int i = 0; i < 5; ++i) {
      const double wsc = W21[j][i] / (double)nn.input_std[i];
      a0 -= wsc * (double)nn.input_mean[i];
      if (i < 3)
        out.hid[j / 2].Ax[i][j % 2] = (float)(kappa * wsc);
      else
        out.hid[j / 2].Au[i - 3][j % 2] = (float)(kappa * wsc);
    }
    out.hid[j / 2].a0[j % 2] = (float)(kappa * a0);
  }
  for (int k = 0; k < 6; ++k) {
    double c0 = nn.b3[k];
    for (int j = 0; j < 32; ++j) {
      c0 += (double)nn.W3[k * 32 + j];
      out.Wo[k][j] = (float)(-2.0 * (double)nn.output_std[k] * (double)nn.W3[k * 32 + j]);
    }
    out.c0[k] = (float)((double)nn.output_std[k] * c0 + (double)nn.output_mean[k]);
  }
}

inline void build_devconst(const AircParams& p, const AircNNBlob& nn, DevConst& d) {
  memset(&d, 0, sizeof(d));
  fuse_network(nn, d.net);
  d.S = (float)p.reference_area;
  d.Sb = (float)(p.reference_area * p.span);
  d.Sc = (float)(p.reference_area * p.chord);
  d.inv_mass = (float)(1.0 / p.mass);
  for (int i = 0; i < 3; ++i) {
    d.com[i] = (float)p.com[i];
    d.g[i] = (float)p.gravity[i];
  }
  for (int i = 0; i < 9; ++i) {
    d.I[i] = (float)p.inertia[i];
    d.Iinv[i] = (float)p.inertia_inv[i];
  }
  d.eps = (float)p.epsilon;
  d.k_rudder = (float)(-0.1 * 6.0 * M_PI / 180.0);
  d.k_flap_x = -0.1f;
  d.k_flap_z = -0.6f;
  d.stall_scaling = p.stall_scaling;
}

// DevConst for one of the registry's other coefficient models (airc_alt.cuh): the network's storage
// holds the model's parameter table.  model_params as documented at airc_create_model.
inline void build_devconst_model(const AircParams& p, int model, const float* model_params, DevConst& d) {
  AircNNBlob* dummy = new AircNNBlob;
  memset(dummy, 0, sizeof(*dummy));
  for (int i = 0; i < 5; ++i) dummy->input_std[i] = 1.0f;
  for (int i = 0; i < 6; ++i) dummy->output_std[i] = 1.0f;
  build_devconst(p, *dummy, d);
  delete dummy;
  AltParams ap;
  memset(&ap, 0, sizeof(ap));
  ap.span = (float)p.span;
  ap.k_rudder_model = (float)((model == AIRC_MODEL_POLY_ ? 0.01 : -0.1) * 6.0 * M_PI / 180.0);
  if (model == AIRC_MODEL_POLY_) {
    memcpy(ap.poly, model_params, sizeof(ap.poly));
    ap.arm = model_params[6 * (NPOLY + 1)];
  } else if (model == AIRC_MODEL_LINEAR_) {
    memcpy(ap.lin, model_params, sizeof(ap.lin));
  }
  memset(&d.net, 0, sizeof(d.net));
  memcpy(&d.net, &ap, sizeof(ap));
  d.model = model;
}

}  // namespace airc
