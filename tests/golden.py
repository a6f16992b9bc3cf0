"""Numbers measured with the reference's own code (deterministic: every sklearn model there uses
random_state=42 and the split uses random_state=42), used as parity fixtures."""
import numpy as np

# hyperparameters_tuning.py, np=4: pooled train accuracy of the pre-averaging local models; rows = hidden sizes in the
# script's order (H:73), columns = learning rates SORTED ascending (0.002 0.004 0.005 0.008 0.01 0.02 0.05 0.1 0.2)
GRID_HIDDEN = [(50,), (100,), (50, 50), (100, 50), (50, 100), (50, 200), (50, 400), (100, 400), (400, 200), (200, 400)]
GRID_LRS = [0.002, 0.004, 0.005, 0.008, 0.01, 0.02, 0.05, 0.1, 0.2]
GRID_ACCURACY = np.array([
    [0.839, 0.832, 0.835, 0.843, 0.844, 0.825, 0.821, 0.814, 0.800],
    [0.856, 0.862, 0.854, 0.848, 0.855, 0.828, 0.811, 0.804, 0.781],
    [0.873, 0.863, 0.878, 0.872, 0.882, 0.845, 0.823, 0.803, 0.512],
    [0.879, 0.921, 0.923, 0.904, 0.877, 0.836, 0.816, 0.804, 0.710],
    [0.888, 0.921, 0.937, 0.906, 0.914, 0.834, 0.811, 0.806, 0.509],
    [0.902, 0.914, 0.944, 0.914, 0.916, 0.827, 0.810, 0.711, 0.639],
    [0.942, 0.957, 0.950, 0.931, 0.925, 0.837, 0.795, 0.641, 0.566],
    [0.960, 0.956, 0.958, 0.941, 0.866, 0.821, 0.813, 0.802, 0.577],
    [0.955, 0.938, 0.931, 0.856, 0.874, 0.839, 0.820, 0.714, 0.488],
    [0.953, 0.950, 0.942, 0.918, 0.907, 0.821, 0.804, 0.803, 0.587],
])
GRID_BEST = {"hidden_layer_sizes": (100, 400), "learning_rate": 0.002, "accuracy": 0.9600}

# FL_SkLearn_MLPClassifier_Limitation.py, np=4, as shipped (every round identical)
SKLEARN_POOLED = {"accuracy": 0.9570, "precision": 0.9584, "recall": 0.9570, "f1": 0.9570}
SKLEARN_WEIGHT_STATS = [((14, 50), -0.006429, 0.285385), ((50, 400), -0.013575, 0.143224), ((400, 1), -0.043961, 0.538777),
                        ((50,), 0.012829, 0.209357), ((400,), -0.012312, 0.098786), ((1,), -0.044904, 0.0)]
