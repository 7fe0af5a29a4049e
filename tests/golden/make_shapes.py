"""Variable shape lists of BASELINE configs 3 and 4, in creation order, derived offline from
random-init architectures on the meta device (no weights, no network):
  resnet50   : torchvision.models.resnet50(weights=None)           161 tensors, 25,557,032 elems
  bert_large : transformers.BertModel(BertConfig(hidden 1024, layers 24, heads 16, ffn 4096,
               vocab 30522))                                        391 tensors, 335,141,888 elems
Writes tests/golden/model_shapes.json. Run from the repo root."""
import json
import os

import torch


def shapes(model):
    return [[name, list(p.shape)] for name, p in model.named_parameters()]


def main():
    import torchvision
    import transformers
    with torch.device("meta"):
        rn = torchvision.models.resnet50(weights=None)
        cfg = transformers.BertConfig(hidden_size=1024, num_hidden_layers=24,
                                      num_attention_heads=16, intermediate_size=4096,
                                      vocab_size=30522)
        bert = transformers.BertModel(cfg)
    out = {"resnet50": shapes(rn), "bert_large": shapes(bert)}
    for k, v in out.items():
        n = sum(int(torch.tensor(s).prod()) if s else 1 for _, s in v)
        print(k, len(v), "tensors", n, "elements")
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "model_shapes.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
