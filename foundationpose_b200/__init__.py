"""foundationpose_b200 -- B200-native (sm_100a) render-and-compare pose-hypothesis path of FoundationPose.

Public surface mirrors the reference: FoundationPose, PoseRefinePredictor, ScorePredictor,
mycpp.cluster_poses, the depth utilities and the FoundationPoseEstimator pipeline processor.
"""
__version__ = '0.1.0'
