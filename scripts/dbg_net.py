import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, torch
from conftest import load_golden
from test_gpu_net import build_net, entry_locations, param_locations
name=sys.argv[1]
arrays, meta = load_golden(name)
eng, spec, entries, x, xd, target, td = build_net(meta, "float64")
B=meta["batch"]
eng.train_begin(optimizer=meta["optimizer"] or "sgd", learning_rate=meta["lr"], decay_rate=meta["decay"], l2_lambda=meta["l2"], epochs=3)
eng.train_step(xd, td, diagnostics=True)
loc=entry_locations(spec)
print(meta["expanded_algos"], spec.rmul, spec.kz, spec.kbn, spec.ka)
for i in range(len(entries)):
    l,w=loc[i]
    got=eng.diag_activation(l,w,B); want=arrays[f"act{i}"]
    e=np.abs(got.reshape(want.shape)-want).max()
    g=eng.diag_activation_grad(l,w,B)
    ge=None
    if g is not None and f"actgrad{i}" in arrays:
        ge=np.abs(g.reshape(arrays[f"actgrad{i}"].shape)-arrays[f"actgrad{i}"]).max()
    print(i, meta["expanded_algos"][i], loc[i], "act err", e, "grad err", ge)
for k,(kind,layer) in enumerate(param_locations(spec, entries)):
    got=eng.download_tensor(kind,layer,"diag_grad"); print(k,kind,layer,np.abs(got-arrays[f"grad{k}"]).max(), np.abs(arrays[f"grad{k}"]).max())
out = eng.forward(xd).cpu().numpy()
print("infer batch err", np.abs(out-arrays["infer_out"]).max())
one = eng.forward(xd[:1].contiguous()).cpu().numpy()[0]
print("one", one[:5], "want", arrays["infer_out_1d"][:5], "batchrow0", out[0][:5], "ref batch row0", arrays["infer_out"][0][:5])
