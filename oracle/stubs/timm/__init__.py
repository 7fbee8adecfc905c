"""Stub of timm 0.5.4 (pinned by the reference's install_vipt.sh:63) -- TEST INFRASTRUCTURE ONLY.
Exposes exactly what the reference hot-path files import. Only `Mlp` carries arithmetic:
fc1 -> act_layer() -> drop -> fc2 -> drop, act_layer = nn.GELU (exact erf)."""
