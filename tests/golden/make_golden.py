"""Regenerates the fixtures under tests/golden/ from the reference tree (run in the build
container, where /root/reference exists; the GPU box only sees the committed outputs).

  heart.txt          = photon-api/src/integTest/resources/DriverIntegTest/input/heart.txt (250 rows x 13)
  a9a_1605.libsvm    = first 1605 rows of .../input/a9a  (the "a1a-shaped" slice)
  reference_kats.json = numbers typed into the reference's own tests, with file:line
"""
import json
import os
import shutil

REF = "/root/reference"
INPUT = os.path.join(REF, "photon-api/src/integTest/resources/DriverIntegTest/input")
HERE = os.path.dirname(os.path.abspath(__file__))

shutil.copyfile(os.path.join(INPUT, "heart.txt"), os.path.join(HERE, "heart.txt"))
with open(os.path.join(INPUT, "a9a")) as f, open(os.path.join(HERE, "a9a_1605.libsvm"), "w") as g:
    for i, line in enumerate(f):
        if i >= 1605:
            break
        g.write(line)

kats = {
    "log1pExp": {
        "source": "photon-lib/src/test/scala/com/linkedin/photon/ml/util/MathUtilsTest.scala:29-34",
        "tolerance": 1e-12,
        "cases": [[-1, 0.313261687518], [0, 0.69314718056], [1, 1.313261687518], [10.5, 10.500027536070],
                  [100.5, 100.5], [10000, 10000]],
    },
    "logistic_identities": {
        "source": "photon-api/src/test/scala/com/linkedin/photon/ml/function/glm/LogisticLossFunctionTest.scala:29-93",
        "tolerance": 1e-12,
        "testCalculate": {"features": [12.21, 10.0, -0.03, 10.3], "coefficients": [1.0, 12.3, -21.0, 0.0],
                          "offset": 1.5},
        "testGradient": {"features": [12.21, 10.0, -0.03, 10.3], "coefficients": [1.0, 1.0, 1.0, 1.0],
                         "offset": 0.0},
        "testHessianVector": [{"features": [0.0, 0.0, 0.0, 0.0], "coefficients": [0.0, 0.0, 0.0, 0.0]},
                              {"features": [1.0, 0.0, 0.0, 0.0], "coefficients": [1.0, 0.0, 0.0, 0.0]}],
    },
    "trivial_labeled_points": {
        "source": "photon-api/src/main/scala/com/linkedin/photon/ml/util/GameTestUtils.scala:70-81",
        "rows": [[0.0, -0.7306653538519616, 0.0], [1.0, 0.6750417712898752, -0.4232874171873786],
                 [1.0, 0.1863463229359709, -0.8163423997075965], [0.0, -0.6719842051493347, 0.0],
                 [1.0, 0.9699938346531928, 0.0], [1.0, 0.22759406190283604, 0.0], [1.0, 0.9688721028330911, 0.0],
                 [0.0, 0.5993795346650845, 0.0], [0.0, 0.9219423508390701, -0.8972778242305388],
                 [0.0, 0.7006904841584055, -0.5607635619919824]],
        "note": "each row = label, x0, x1; the tests append an intercept column of 1.0",
    },
    "ridge_golden": {
        "source": "photon-api/src/integTest/scala/com/linkedin/photon/ml/estimators/GameEstimatorIntegTest.scala:61-105",
        "task": "LINEAR_REGRESSION", "l2": 0.3, "tolerance": 1e-12,
        "coefficients": [0.3215554473500486, 0.17904355431985355, 0.4122241763914806],
    },
    "normalization_golden": {
        "source": "photon-api/src/integTest/scala/com/linkedin/photon/ml/estimators/GameEstimatorIntegTest.scala:129-179",
        "task": "LINEAR_REGRESSION", "l2": 0.0, "tolerance": 1e-8,
        "coefficients": [0.34945501725815586, 0.26339479490270173, 0.4366125400310442],
        "normalization_types": ["NONE", "SCALE_WITH_MAX_MAGNITUDE", "SCALE_WITH_STANDARD_DEVIATION", "STANDARDIZATION"],
    },
}
with open(os.path.join(HERE, "reference_kats.json"), "w") as f:
    json.dump(kats, f, indent=1)
print("wrote fixtures to", HERE)
