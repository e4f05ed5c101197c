"""`keypoint_moseq.gibbs` of the reference, served by keypointmoseq_b200.gibbs (ref: gibbs.py:10-156)."""
from keypointmoseq_b200.gibbs import *  # noqa: F401,F403
from keypointmoseq_b200.gibbs import (resample_ar_params, resample_hdp_transitions, resample_heading,  # noqa: F401
                                      resample_latents, resample_location, resample_obs_params,
                                      resample_obs_variance, resample_regression_params, resample_scales,
                                      resample_stateseqs, resample_transitions)
