# BASELINE.json config 4 -- ViT-S/16 sizes (E=384, depth 12, heads 6, MLP 1536) with the FFT mixer,
# 224^2 synthetic ImageNet-shaped input -> 14x14 token grid, bf16, batch 256 per GPU
__base__ = "default.py"

mixer = "fft"
precision = "bf16"
batch_size = 256
bench_batch_size = 2048         # per-GPU batch of bench.py: sized for 180 GB HBM3e (56 GB peak); a mixer launch costs ~12 us + 2.6 / 4.2 us per tile (tools/tc_fixed_cost.py)
val_batch_size = 256
epochs = 300
num_classes = 1000
patch_size = 16
img_size = 224
in_channels = 3
num_heads = 6
dropout = 0.001
hidden_dim = 1536
adam_weight_decay = 0.01
adam_betas = (0.9, 0.999)
activation = "gelu"
num_encoders = 12
embed_dim = 384
num_patches = (img_size // patch_size) ** 2
