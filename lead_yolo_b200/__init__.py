"""lead_yolo_b200 — B200-native (sm_100a) kernels for the LEAD-YOLO module hot path."""
__version__ = "0.1.0"
