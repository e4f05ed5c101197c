"""`keypoint_moseq.kalman` of the reference (ref: kalman.py:80-156): the generic one-sequence entry
points, served by the associative-scan variant.  The Gibbs steps do not go through these — `resample_latents`
and `resample_location` have their own kernels for the structured models they build."""
from keypointmoseq_b200.kalman_scan import kalman_filter, kalman_sample  # noqa: F401
