"""`smoothquant.calibration` import path (smoothquant/calibration.py:13-51) -> GPU collector (mq_colstat, no CPU
fallback).  Same signature as the reference; the dataset / tokenizer plumbing is the reference's, the per-linear
statistic runs in K0.  ``get_static_decoder_layer_scales`` (INT8 OPT export, :54-130) is out of scope."""
from smoothquant_mixedprecision_b200 import calibration as _gpu


def get_act_scales(model, tokenizer, dataset_path, num_samples=512, seq_len=512):
    from datasets import load_dataset                       # not installed on the build box: imported on use
    dataset = load_dataset("json", data_files=dataset_path, split="train").shuffle(seed=42)
    batches = (tokenizer(dataset[i]["text"], return_tensors="pt", max_length=seq_len, truncation=True).input_ids
               for i in range(num_samples))
    return _gpu.get_act_scales(model, batches)
