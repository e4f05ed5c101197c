// Compile-time (latent dim D, lags L) combinations the register-tiled kernels are built
// for.  Adding a pair is one line here.  The first entries are the BASELINE.json
// configurations (D=10,L=3 and D=7,L=3); the small ones serve brute-force-checkable tests.
#pragma once

#define KPMS_FOR_EACH_DL(X) \
  X(10, 3)                  \
  X(7, 3)                   \
  X(10, 2)                  \
  X(10, 1)                  \
  X(5, 3)                   \
  X(4, 2)                   \
  X(3, 2)                   \
  X(2, 2)

static inline bool kpms_dl_supported(int D, int L) {
#define KPMS_X(d, l) \
  if (D == d && L == l) return true;
  KPMS_FOR_EACH_DL(KPMS_X)
#undef KPMS_X
  return false;
}
