"""Detections wire format of the reference (src/dataset.py:97-109 exportDetections, :133-141
createDetectionsFromPath):

    {"views": [{"sensorPoints": [[u, v], ...], "modelPoints": [[X, Y, Z], ...]}, ...]}

so that detections written by the reference (or by any detector that follows it) reach the GPU path,
and back.  Files written here are byte-identical to the reference's for the same detections."""
import json

import numpy as np


def exportDetections(allDetections, filePath):
    """allDetections: list of (sensorPoints (N, 2), modelPoints (N, 3)) per view."""
    detectionsDict = {"views": [{"sensorPoints": np.asarray(s).tolist(), "modelPoints": np.asarray(m).tolist()}
                                for s, m in allDetections]}
    with open(filePath, "w") as f:
        f.write(json.dumps(detectionsDict))


def createDetectionsFromPath(filePath):
    with open(filePath, "r") as f:
        detectionsDict = json.load(f)
    return [(np.array(view["sensorPoints"]).reshape(-1, 2), np.array(view["modelPoints"]).reshape(-1, 3))
            for view in detectionsDict["views"]]
