"""TEST INFRASTRUCTURE: the CPU oracle (oracle/mq_oracle.py) wrapped as nn.Modules, so that whole models can be run
through the reference's arithmetic on the host CPU of the GPU box and compared with the CUDA path layer by layer.

Nothing here is imported by the product package.  The walkers below mirror which linears the reference swaps
(fake_quant.py:377-461 OPT, :464-561 Llama) and with which `quantize_output` flag.
"""
import functools

import torch
from torch import nn

from oracle import mq_oracle as O


class OracleRefLinear(nn.Module):
    """Literal reference layer: W4A4Linear.from_float + forward (fake_quant.py:279-371) through the oracle."""

    def __init__(self, lin, importance, salient_prop, group_size, quantize_output, weight_quant="per_group",
                 act_quant="per_group"):
        super().__init__()
        self.salient = O.select_salient(None if importance is None else importance.float(), salient_prop)
        self.g, self.qo, self.aq = group_size, quantize_output, act_quant
        self.weight_q = O.w4a4_weight_reference(lin.weight.data, self.salient, 4, group_size, weight_quant)
        self.bias = None if lin.bias is None else lin.bias.data.clone()
        self.in_features, self.out_features = lin.in_features, lin.out_features

    def forward(self, x, quantize_output=None):
        qo = self.qo if quantize_output is None else quantize_output
        return O.w4a4_forward_reference(x, self.weight_q, self.bias, self.salient, 4, self.g, act_quant=self.aq,
                                        quantize_output=qo)


class OracleFusedLinear(nn.Module):
    """mode="fused" contract (SURVEY 8(a')): one static partition, a2(x[:, perm]) @ a4(W[:, perm])^T + salient + b,
    evaluated from the oracle's codes and scales; the output quantiser (if any) is the literal dynamic-sort one."""

    def __init__(self, lin, importance, salient_prop, group_size, quantize_output, sort_key=None):
        super().__init__()
        K = lin.in_features
        self.salient = O.select_salient(None if importance is None else importance.float(), salient_prop)
        key = sort_key if sort_key is not None else importance
        self.part = O.fused_partition(K, self.salient, None if key is None else key.float(), group_size)
        self.g, self.qo = group_size, quantize_output
        w = lin.weight.data
        self.cdt = w.dtype if w.dtype in (torch.float16, torch.bfloat16) else torch.float16
        self.w = w.to(self.cdt)
        self.wq = O.group_quant(self.w[:, self.part.perm], 4, group_size)
        self.w_val = _exact_values(self.wq)
        self.bias = None if lin.bias is None else lin.bias.data.to(self.cdt)
        self.in_features, self.out_features = K, lin.out_features

    def pre_quant_f32(self, x2):
        """fp32 evaluation of the composed oracle (codes are small integers, so each group's partial sum is exact in
        fp32; the scale products and the sum over groups round at 2^-24 -- far below the fp16 tolerance)."""
        xc = x2.to(self.cdt)
        aq = O.group_quant(xc[:, self.part.perm], 4, self.g)
        y = _exact_values(aq) @ self.w_val.t()
        if self.part.salient.numel():
            y += xc[:, self.part.salient].float() @ self.w[:, self.part.salient].float().t()
        if self.bias is not None:
            y += self.bias.float().view(1, -1)
        return y

    def pre_quant(self, x2):
        return self.pre_quant_f32(x2).to(self.cdt)

    def forward(self, x, quantize_output=None):
        shape = x.shape
        y = self.pre_quant(x.reshape(-1, shape[-1]))
        if self.qo if quantize_output is None else quantize_output:
            y = output_quant_oracle(y, self.salient, self.g)
        return y.to(x.dtype).view(*shape[:-1], -1)


def _exact_values(q):
    """code * scale as fp32 (3-bit code x 11-bit scale: exact), [R, G*g]"""
    R, G = q.scales.shape
    return (q.codes.view(R, G, q.group_size).float() * q.scales.float()[:, :, None]).view(R, G * q.group_size)


def output_quant_oracle(y, salient, group_size, n_bits=4):
    """fake_quant.py:309-316: the act quantiser applied to y with the K-sized salient mask on y's columns."""
    if salient is None:
        return O.act_quant_group_sort(y, n_bits, group_size)[0]
    keep = O.nonsalient_columns(y.shape[-1], salient)
    out = y.clone()
    out[:, keep] = O.act_quant_group_sort(y[:, keep], n_bits, group_size)[0]
    return out


OPT_LINEARS = {"fc1": False, "fc2": False, "q_proj": True, "k_proj": True, "v_proj": True, "out_proj": False}
LLAMA_LINEARS = {"gate_proj": False, "up_proj": False, "down_proj": False, "q_proj": False, "k_proj": False,
                 "v_proj": False, "o_proj": False}


def swapped_linears(model, kind):
    """(parent module, attribute, full name, quantize_output) of every linear the reference walkers replace."""
    table = OPT_LINEARS if kind == "opt" else LLAMA_LINEARS
    out = []
    for name, m in model.model.named_modules():
        for attr, qo in table.items():
            child = getattr(m, attr, None)
            if isinstance(child, nn.Module) and hasattr(child, "in_features"):
                out.append((m, attr, "model." + name + "." + attr, qo))
    return out


def oracle_swap(model, kind, mode, input_feat, salient_prop, group_size, sort_keys=None):
    """Replace the walker's linears by oracle-backed modules (CPU).  importance = sum of calibration blocks
    (fake_quant.py:396, :486)."""
    for parent, attr, full, qo in swapped_linears(model, kind):
        lin = getattr(parent, attr)
        imp = sum(input_feat[full]).float()
        if mode == "reference":
            new = OracleRefLinear(lin, imp, salient_prop, group_size, qo)
        else:
            new = OracleFusedLinear(lin, imp, salient_prop, group_size, qo,
                                    None if sort_keys is None else sort_keys[full])
        setattr(parent, attr, new)
    return model


def hook_input_feat(model, ids_list):
    """run_experiments.get_calib_feat (run_experiments.py:55-84): per nn.Linear, per block, mean|x| over tokens."""
    feat, hooks = {}, []

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        feat.setdefault(name, []).append(x.reshape(-1, x.shape[-1]).abs().mean(dim=0).cpu().detach())

    for name, m in model.named_modules():
        if isinstance(m, nn.Linear):
            hooks.append(m.register_forward_hook(functools.partial(hook, name=name)))
    with torch.no_grad():
        for ids in ids_list:
            model(ids)
    for h in hooks:
        h.remove()
    return feat


def hook_act_scales(model, ids_list):
    """calibration.get_act_scales (calibration.py:18-25): running per-channel max|x|."""
    scales, hooks = {}, []

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        cur = x.reshape(-1, x.shape[-1]).abs().detach().max(dim=0)[0].float().cpu()
        scales[name] = torch.max(scales[name], cur) if name in scales else cur

    for name, m in model.named_modules():
        if isinstance(m, nn.Linear):
            hooks.append(m.register_forward_hook(functools.partial(hook, name=name)))
    with torch.no_grad():
        for ids in ids_list:
            model(ids)
    for h in hooks:
        h.remove()
    return scales


def capture_io(model, classes):
    """Forward hooks recording (input, output) of every module of the given classes, by name.  Returns (dict, remove)."""
    rec, hooks = {}, []

    def hook(m, x, y, name):
        x = x[0] if isinstance(x, tuple) else x
        rec[name] = (x.detach().clone(), y.detach().clone())

    for name, m in model.named_modules():
        if isinstance(m, classes):
            hooks.append(m.register_forward_hook(functools.partial(hook, name=name)))

    def remove():
        for h in hooks:
            h.remove()
    return rec, remove


def margin_top1_agree(logits, ref):
    """Arg-max agreement wherever the reference's top-1 margin exceeds 4x the largest logit difference."""
    top2 = ref.topk(2, dim=-1).values
    sure = (top2[..., 0] - top2[..., 1]) > 4 * (logits - ref).abs().max()
    return bool(torch.equal(logits.argmax(-1)[sure], ref.argmax(-1)[sure])), int(sure.sum())
