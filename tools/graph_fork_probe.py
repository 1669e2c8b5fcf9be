#!/usr/bin/env python
"""Which op families keep two captured branches from running concurrently?  (CUDA-graph fork/join probe)"""
import torch

w = torch.randn(512, 512, device="cuda")
lstm = torch.nn.LSTM(512, 512, 1, bias=False).cuda()
lstm.flatten_parameters()
emb = torch.nn.Embedding(1000, 512).cuda()
idx = torch.randint(0, 1000, (6,), device="cuda")


def chain_mm(t, n=300):
    for _ in range(n):
        t = torch.tanh(t @ w * 0.01) + 1.0
    return t


def chain_lstm(t, n=100):
    st = (torch.zeros(1, 6, 512, device="cuda"), torch.zeros(1, 6, 512, device="cuda"))
    for _ in range(n):
        o, st = lstm(t.unsqueeze(0), st)
        t = o.squeeze(0)
    return t


def chain_softmax(t, n=300):
    for _ in range(n):
        t = torch.log_softmax(t, dim=1) + emb(idx)
    return t


s1 = torch.cuda.Stream()
for name, chain in (("matmul+tanh", chain_mm), ("cudnn lstm", chain_lstm), ("log_softmax+embedding", chain_softmax)):
    x = [torch.randn(6, 512, device="cuda") for _ in range(2)]
    res = {}
    for par in (False, True):
        g = torch.cuda.CUDAGraph()
        with torch.no_grad():
            chain(x[0], 3)
            with torch.cuda.stream(s1):
                chain(x[1], 3)
            torch.cuda.synchronize()
            with torch.cuda.graph(g):
                main = torch.cuda.current_stream()
                if par:
                    s1.wait_stream(main)
                    with torch.cuda.stream(s1):
                        b = chain(x[1])
                    a = chain(x[0])
                    main.wait_stream(s1)
                else:
                    a = chain(x[0])
                    b = chain(x[1])
                c = a + b
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        res[par] = e0.elapsed_time(e1) / 5
    print(f"{name:24s} serial {res[False]:.2f} ms   forked {res[True]:.2f} ms")
