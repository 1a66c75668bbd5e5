"""Logging / path / model-I/O helpers the four pruning modules import, with the reference's names
and CSV wire formats (/root/reference/ResNet/utils.py): the per-layer log (9 columns, :55-58) is read
back by min_weight_pruning and random_pruning by column name, the experiment log has 8 columns
(:127-128).  I/O only -- nothing here is on the GPU path."""
import csv
import os

PER_LAYER_FIELDS = ['GUID', 'Wavelet', 'Level', 'Threshold', 'DWT Phase', 'Original Parameter Count', 'Non-zero Params',
                    'Total Pruned Count', 'Layer Name']
EXPERIMENT_FIELDS = ['GUID', 'Wavelet', 'Level', 'Threshold', 'Phase', 'Total Pruned Count', 'Total Non-Zero Params',
                     'Model Path']

# the reference hard-codes ../../WaveletTransforms/ResNet/SavedModels (utils.py:148-162); the
# directory is configurable here and defaults to ./SavedModels
OUTPUT_ROOT = os.environ.get('WT_OUTPUT_DIR', 'SavedModels')


def load_model(model_path, config_path=None):
    """utils.py:6-25 -- HF AutoModelForImageClassification on the CPU, configuration read from `config_path`
    (the model directory itself when omitted); ValueError when `model_path` is not a directory (:17-22)"""
    from transformers import AutoConfig, AutoModelForImageClassification
    if not os.path.isdir(model_path):
        raise ValueError(f"Provided model path {model_path} is not a valid directory.")
    config = AutoConfig.from_pretrained(config_path if config_path else model_path)
    model = AutoModelForImageClassification.from_pretrained(model_path, config=config)
    print("Pre-trained model loaded successfully.")
    return model


_save_pool = None
_pending_saves = []


def _write_model(model, output_dir, state_dict=None):
    os.makedirs(output_dir, exist_ok=True)
    if hasattr(model, 'save_pretrained'):
        if state_dict is None:
            model.save_pretrained(output_dir)
        else:
            model.save_pretrained(output_dir, state_dict=state_dict)
    else:
        import torch
        torch.save(model.state_dict() if state_dict is None else state_dict, os.path.join(output_dir, 'pytorch_model.bin'))


def save_model(model, output_dir):
    """utils.py:28-38.  With config.async_save the tensors are snapshotted here (device tensors through pinned
    host memory, non-blocking) and the serialisation runs on a background thread; wait_for_saves() joins."""
    from . import config
    if not config.async_save:
        _write_model(model, output_dir)
        return None
    import atexit
    from concurrent.futures import ThreadPoolExecutor

    import torch
    global _save_pool
    if _save_pool is None:
        _save_pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix='wt-save')
        atexit.register(wait_for_saves)
    snap, on_device = {}, False
    for k, v in model.state_dict().items():
        v = v.detach()
        if v.is_cuda:
            h = torch.empty(v.shape, dtype=v.dtype, pin_memory=True)
            h.copy_(v, non_blocking=True)
            snap[k], on_device = h, True
        else:
            snap[k] = v.clone()
    if on_device:
        torch.cuda.current_stream().synchronize()
    fut = _save_pool.submit(_write_model, model, output_dir, snap)
    _pending_saves.append(fut)
    return fut


def wait_for_saves():
    """joins every background save started by save_model; re-raises the first failure"""
    while _pending_saves:
        _pending_saves.pop(0).result()


def setup_csv_writer(file_path, mode='w'):
    """utils.py:41-65"""
    file_exists = os.path.isfile(file_path)
    d = os.path.dirname(file_path)
    if d:
        os.makedirs(d, exist_ok=True)
    f = open(file_path, mode=mode, newline='')
    writer = csv.DictWriter(f, fieldnames=PER_LAYER_FIELDS)
    if mode == 'w' or (mode == 'a' and not file_exists):
        writer.writeheader()
    return writer, f


def log_pruning_details(csv_writer, guid, wavelet, level, threshold, phase, original_param_count, non_zero_params,
                        total_pruned_count, layer_name):
    """utils.py:68-101"""
    if not csv_writer or not layer_name:
        return
    csv_writer.writerow({
        'GUID': guid, 'Wavelet': wavelet, 'Level': level, 'Threshold': threshold, 'DWT Phase': phase,
        'Original Parameter Count': original_param_count, 'Non-zero Params': non_zero_params,
        'Total Pruned Count': total_pruned_count, 'Layer Name': layer_name})


def append_to_experiment_log(file_path, guid, wavelet, level, threshold, phase, total_pruned_count, total_non_zero_params,
                             model_path):
    """utils.py:104-145"""
    if not file_path:
        print('Error: Invalid file path provided.')
        return
    file_path = os.path.normpath(file_path)
    file_exists = os.path.isfile(file_path)
    d = os.path.dirname(file_path)
    if d:
        os.makedirs(d, exist_ok=True)
    with open(file_path, mode='a', newline='') as f:
        writer = csv.DictWriter(f, fieldnames=EXPERIMENT_FIELDS)
        if not file_exists:
            writer.writeheader()
        writer.writerow({
            'GUID': guid, 'Wavelet': wavelet, 'Level': level, 'Threshold': threshold, 'Phase': phase,
            'Total Pruned Count': total_pruned_count, 'Total Non-Zero Params': total_non_zero_params,
            'Model Path': model_path})


def check_and_set_pruned_instance_path(pruned_instance):
    """utils.py:148-162"""
    path = os.path.join(OUTPUT_ROOT, pruned_instance)
    os.makedirs(path, exist_ok=True)
    return path


def print_model_summary(model):
    """utils.py:165-190 -- weight shapes and parameter counts of the model's direct children"""
    total_params = 0
    print("Model Summary:")
    print("Layer Name" + "\t" * 7 + "Output Shape" + "\t" * 5 + "Param #")
    print("=" * 100)
    for name, module in model.named_children():
        if hasattr(module, 'weight') and hasattr(module.weight, 'size'):
            print(f"{name}\t{module.weight.size()}\t{module.weight.numel()}")
            total_params += module.weight.numel()
            if hasattr(module, 'bias') and hasattr(module.bias, 'size'):
                total_params += module.bias.numel()
    print("=" * 100)
    print(f"Total Params: {total_params}")


def print_model_structure(model, depth=0):
    """utils.py:193-206 -- the module tree, two spaces per level"""
    indent = " " * (depth * 2)
    for name, module in model.named_children():
        print(f"{indent}{name} - {module.__class__.__name__}")
        if list(module.children()):
            print_model_structure(module, depth + 1)


def get_layer(model, layer_name):
    """utils.py:209-238 -- dotted lookup; None when a component is missing; a leading
    'ResNetForImageClassification.' (:220-223) and empty components are ignored"""
    prefix = 'ResNetForImageClassification.'
    if layer_name.startswith(prefix):
        layer_name = layer_name[len(prefix):]
    cur = model
    for part in layer_name.split('.'):
        if not part:
            continue
        if part.isdigit() and hasattr(cur, '__getitem__'):
            try:
                cur = cur[int(part)]
                continue
            except (IndexError, KeyError, TypeError):
                return None
        if not hasattr(cur, part):
            return None
        cur = getattr(cur, part)
    return cur
