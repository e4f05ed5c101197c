"""`keypoint_moseq.util` of the reference, served by keypointmoseq_b200.util (ref: util.py:12-275)."""
from keypointmoseq_b200.util import *  # noqa: F401,F403
from keypointmoseq_b200.util import (affine_transform, angle_to_rotation_matrix, ar_to_lds, center_embedding,  # noqa: F401
                                     count_transitions, ensure_symmetric, expected_keypoints, get_lags,
                                     interpolate_keypoints, inverse_affine_transform, merge_data, pad_affine,
                                     pca_transform, stateseq_stats, vector_to_angle)
from keypointmoseq_b200.gibbs import log_joint_likelihood  # noqa: F401  (ref: util.py:266-275)
