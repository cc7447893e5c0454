# BASELINE.json config 1 -- hyper-parameters of the reference's spectre_vit/configs/spectre_vit_cifar100.py:3-20
__base__ = "default.py"

mixer = "fft"
precision = "fp32"
batch_size = 8
val_batch_size = 512
epochs = 1000
num_classes = 100
patch_size = 4
img_size = 32
in_channels = 3
num_heads = 16
dropout = 0.001
hidden_dim = 768
adam_weight_decay = 0.01
adam_betas = (0.9, 0.999)
activation = "gelu"
num_encoders = 4
embed_dim = 512
num_patches = (img_size // patch_size) ** 2
