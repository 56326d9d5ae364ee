import sys, numpy as np, torch
sys.path.insert(0, ".")
from kafka_inferenceengine_b200 import synthetic as syn
from kafka_inferenceengine_b200.device import DeviceVector
from kafka_inferenceengine_b200.inference.utils import emulator_table, linearize_bands
dev = torch.device("cuda:0")
rng = np.random.default_rng(21)
P = 10
gp = syn.make_emulator(rng, syn.SAIL_MEAN, syn.SAIL_SIGMA, n_train=250, cls=syn.EmulatorData)
for n in (1000, 100000, 1883756):
    gen = torch.Generator(device=dev).manual_seed(5)
    x = DeviceVector.empty(n, P, dev)
    x.data.copy_(torch.from_numpy(syn.SAIL_MEAN).to(dev)[:, None] + torch.from_numpy(syn.SAIL_SIGMA).to(dev)[:, None] *
                 torch.randn((P, x.ld), generator=gen, dtype=torch.float64, device=dev))
    tab = emulator_table(gp, dev)
    h0, jac = linearize_bands(x, [(tab, np.arange(P))], None, False)
    h0b, jacb = linearize_bands(x, [(tab, np.arange(P))], None, False)
    print(n, "repeat equal:", torch.equal(h0[:, :n], h0b[:, :n]), torch.equal(jac[:, :, :n], jacb[:, :, :n]))
    perm = torch.randperm(n, generator=gen, device=dev)
    xp = DeviceVector.empty(n, P, dev)
    xp.data.zero_()
    xp.data[:, :n] = x.data[:, :n][:, perm]
    h0p, jacp = linearize_bands(xp, [(tab, np.arange(P))], None, False)
    a, b = h0p[0, :n], h0[0, :n][perm]
    bad = (a != b).nonzero().flatten()
    print("  mismatches", bad.numel(), "max rel", ((a - b).abs() / b.abs()).max().item())
    if bad.numel():
        i = bad[:10].cpu().numpy(); print("  new idx", i, "slot", (i % 256) // 128, "old idx", perm[bad[:10]].cpu().numpy(), "slot", (perm[bad[:10]].cpu().numpy() % 256) // 128)
