# BASELINE.json config 2 -- hyper-parameters of the reference's spectre_vit/configs/spectre_vit_mnist.py:3-19
__base__ = "default.py"

mixer = "fft"
precision = "fp32"
batch_size = 8
bench_batch_size = 4096          # bandwidth-measurable batch
val_batch_size = 512
epochs = 1000
num_classes = 100
patch_size = 4
img_size = 28
in_channels = 3
num_heads = 8
dropout = 0.001
hidden_dim = 256
adam_weight_decay = 0.01
adam_betas = (0.9, 0.999)
activation = "gelu"
num_encoders = 4
embed_dim = (patch_size**2) * in_channels   # 48
num_patches = (img_size // patch_size) ** 2  # 49 -> 7x7 grid
