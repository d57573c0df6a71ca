"""Write the 'pretrained' file Stoke-DDP.py:209-213 loads (random-init weights: there is no network for the real one)."""
import os

import torch
from pytorch_distributedtraining_b200.models.swinir import SwinIR

torch.manual_seed(0)
m = SwinIR(upscale=2, in_chans=3, img_size=64, window_size=8, img_range=1., depths=[6, 6, 6, 6], embed_dim=60,
           num_heads=[6, 6, 6, 6], mlp_ratio=2, upsampler='pixelshuffledirect', resi_connection='1conv')
os.makedirs("model_zoo", exist_ok=True)
torch.save({"params": m.state_dict()}, "model_zoo/002_lightweightSR_DIV2K_s64w8_SwinIR-S_x2.pth")
print("wrote model_zoo/002_lightweightSR_DIV2K_s64w8_SwinIR-S_x2.pth")
