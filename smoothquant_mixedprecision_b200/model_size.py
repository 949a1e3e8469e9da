"""`get_model_size` under the reference's name and signature (smoothquant/model_size.py:5-15), returning BITS like the
reference (its caller divides by ``MiB = 1024 * 1024 * 8``, run_experiments.py:24-28, :178-179).

The reference's formula is analytic: ``numel(model.parameters()) * average data width``, with ``(16 + 4) /
group_size`` bits of scale overhead per element.  After its own walkers ran, the quantised linears are *buffers*, not
parameters, so it stops counting them.  Here every swapped ``W4A4Linear`` is counted with the bytes it
really holds at rest (int4 nibbles + 16-bit group scales + 16-bit salient columns + permutation, ``packed_bytes``);
everything that is not a swapped linear is counted with the reference's formula, so a float model gives the
reference's number exactly."""
from torch import nn


def get_model_size(model: nn.Module, data_width=16, salient_prop=0, group_size=-1):
    from .fake_quant import W4A4Linear
    data_width_non_salient = data_width
    data_width_salient = 16
    if group_size != -1:
        data_width_non_salient += (16 + 4) / group_size
        data_width_salient += (16 + 4) / group_size
    avg_data_width = data_width_non_salient * (1 - salient_prop) + data_width_salient * salient_prop
    num_elements = 0
    for param in model.parameters():
        num_elements += param.numel()
    packed_bits = sum(m.packed_bytes() * 8 for m in model.modules() if isinstance(m, W4A4Linear))
    return num_elements * avg_data_width + packed_bits
