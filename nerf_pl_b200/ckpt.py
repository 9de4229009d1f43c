"""Checkpoint I/O in the reference's Lightning layout (utils/__init__.py:55-76,
utils/save_weights_only.py:12-15).

A Lightning checkpoint is `{'state_dict': {'nerf_coarse.xyz_encoding_1.0.weight': ..., 'nerf_fine....': ...}, ...}`;
`save_weights_only.py` re-saves the bare `state_dict`.  `extract_model_state_dict` / `load_ckpt` keep the reference's
names, arguments and behaviour; loading copies IN PLACE (`load_state_dict`), so parameters that a fused optimizer has
re-homed into its flat buffer stay views of it and their version counters trigger the bf16 operand re-pack on the next
render.  `save_ckpt` writes the layout the reference's loaders (eval.py:104-105) read."""
import torch


def extract_model_state_dict(ckpt_path, model_name='model', prefixes_to_ignore=[]):
    checkpoint = torch.load(ckpt_path, map_location=torch.device('cpu'))
    if 'state_dict' in checkpoint:          # a pytorch-lightning checkpoint
        checkpoint = checkpoint['state_dict']
    picked = {}
    for key, value in checkpoint.items():
        if not key.startswith(model_name):
            continue
        key = key[len(model_name) + 1:]
        if any(key.startswith(prefix) for prefix in prefixes_to_ignore):
            print('ignore', key)
            continue
        picked[key] = value
    return picked


def load_ckpt(model, ckpt_path, model_name='model', prefixes_to_ignore=[]):
    model_dict = model.state_dict()
    model_dict.update(extract_model_state_dict(ckpt_path, model_name, prefixes_to_ignore))
    model.load_state_dict(model_dict)


def save_ckpt(ckpt_path, models, lightning=True, **extra):
    """models: {'nerf_coarse': module, 'nerf_fine': module}.  lightning=True nests the tensors under 'state_dict' (what
    Lightning's ModelCheckpoint writes, train.py:154-158); False writes the bare dict of save_weights_only.py."""
    state = {f"{name}.{k}": v.detach().cpu().clone() for name, m in models.items() for k, v in m.state_dict().items()}
    torch.save(dict(state_dict=state, **extra) if lightning else state, ckpt_path)
