"""Host-side cost of the stream primitives a pipelined step needs (microseconds per call, wall clock)."""
import time
import torch

dev = torch.device("cuda", 0)
x = torch.zeros(1024, device=dev)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    x.add_(1.0)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
ev = torch.cuda.Event()
user = torch.cuda.current_stream()


def t(name, fn, n=2000):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    dt = (time.perf_counter() - t0) / n * 1e6
    torch.cuda.synchronize()
    print(f"{name:40s} {dt:7.2f} us")


t("graph.replay()", g.replay)
t("event.record(stream)", lambda: ev.record(s1))
t("stream.wait_event(ev)", lambda: s2.wait_event(ev))


def ctx():
    with torch.cuda.stream(s1):
        pass


t("with torch.cuda.stream(s): pass", ctx)


def setstream():
    torch.cuda.set_stream(s1)
    torch.cuda.set_stream(user)


t("set_stream(s); set_stream(user)", setstream)


def replay_on():
    torch.cuda.set_stream(s1)
    g.replay()
    torch.cuda.set_stream(user)


t("set_stream + replay + set_stream", replay_on)
t("torch.cuda.current_stream()", lambda: torch.cuda.current_stream(dev))
t("user.wait_stream(s1)", lambda: user.wait_stream(s1))
