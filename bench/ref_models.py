"""Models handed to the UNMODIFIED reference Trainer for the driver-named workloads (BASELINE.json configs 3-4).

The reference ships only MLModel; ResNet-50 and BERT-base plug into its "bring your own nn.Module" slot
(reference main.py:30).  These are the stock library definitions (torchvision / HuggingFace), i.e. what a
user of the reference would pass.  bf16 legs: the reference has no AMP, so the model is cast up front and the
wrapper casts floating inputs.  Dropout is set to 0 in both arms.
"""
import torch
from torch import nn


class _HFLogits(nn.Module):
    """The reference calls ``model(inputs)`` and feeds the result to CrossEntropyLoss: return plain fp32 logits."""

    def __init__(self, hf):
        super().__init__()
        self.hf = hf

    def forward(self, input_ids):
        return self.hf(input_ids=input_ids).logits.float()


class _CastInput(nn.Module):
    def __init__(self, m, dtype):
        super().__init__()
        self.m, self.dtype = m, dtype

    def forward(self, x):
        return self.m(x.to(self.dtype)).float()


def build_reference_model(name: str, dtype: str):
    td = torch.bfloat16 if dtype == "bf16" else torch.float32
    if name == "bert_base":
        import transformers

        cfg = transformers.BertConfig(num_labels=2, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
        return _HFLogits(transformers.BertForSequenceClassification(cfg).to(td))
    if name == "resnet50":
        import torchvision

        return _CastInput(torchvision.models.resnet50(weights=None).to(td), td)
    raise KeyError(name)
