from .kitti import (KittiOdometry, KittiOdometryVelodyneData, KittiOdometryVelodyneSequenceData, cam2velo, velo2cam,
                    read_velodyne_bin, stage_scans)

__all__ = ['KittiOdometry', 'KittiOdometryVelodyneData', 'KittiOdometryVelodyneSequenceData', 'cam2velo', 'velo2cam',
           'read_velodyne_bin', 'stage_scans']
