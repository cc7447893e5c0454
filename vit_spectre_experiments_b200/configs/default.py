random_seed = 42
learning_rate = 1e-3
