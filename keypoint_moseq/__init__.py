"""Drop-in package name of the reference (calebweinreb/keypointMoSeq): `keypoint_moseq.gibbs`,
`keypoint_moseq.initialize`, `keypoint_moseq.util`, `keypoint_moseq.distributions`,
`keypoint_moseq.transitions` and `keypoint_moseq.kalman` resolve to the B200 implementations in `keypointmoseq_b200`, so the
notebook cells (`from keypoint_moseq.gibbs import *` ...) run with torch CUDA tensors in place of
JAX arrays.  Nothing here computes; there is no JAX, CPU or other fallback behind these names."""
