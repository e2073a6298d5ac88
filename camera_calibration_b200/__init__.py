"""calib-b200: B200-native (sm_100a) Levenberg-Marquardt refinement stage of Zhang's-method
camera calibration, a drop-in for pvphan/camera-calibration's
Calibrator.refineCalibrationParameters (src/calibrate.py:117-171).  See DESIGN.md."""
from . import dataset, distortion, linearcalibrate, mathutils, synth  # noqa: F401
from .calibrate import (BatchProblem, Calibrator, MultiProblem, Problem, ProjectionJacobian, calibrateCamera,  # noqa: F401
                        flatten_detections, getSensorPoints)
from ._lib import CalibError, build  # noqa: F401
