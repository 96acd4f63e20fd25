"""lidar_object_detection_b200 -- B200-native per-point LiDAR-camera fusion and evaluation.

``lidarfuse``   the reference's function surface (V1-V5, cvs_erosion, seg_with_pointcloud)
                on hand-written sm_100a CUDA behind include/lidarfuse.h
``kitti360``    calibration / sweep / box loaders (host numpy float64)
``matching2d``  the V4 / V5 2D matching helpers (host) and their batched device versions
``colour``      headless colour buffers (per-point RGB, depth-map overlay)
``io``          sweep streaming, packed-mask container, batched CSV writer
``compare``     master CSV vs the reference's recorded spreadsheet (xlsx read as zip + XML)
``driver``      the cvs_erosion.process_frames loop -> master_car_statistics.csv
``synthetic``   HDL-64-shaped synthetic batches for the benchmark
``distributed`` frame sharding over torch.distributed + the counter all-reduce
"""
from . import _abi  # noqa: F401

__all__ = ['build', 'lidarfuse', 'kitti360', 'matching2d', 'colour', 'io', 'compare', 'driver', 'synthetic', 'distributed']
