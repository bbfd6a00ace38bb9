import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omldm_b200.engine import Engine
from omldm_b200.stream import ring as R
from omldm_b200.utils.config import JobParameters
dev = torch.device("cuda", 0)
d, B = 28, 16384
p = JobParameters(parallelism=1); p.extra["graphs"] = "false"
eng = Engine(p, 0, 1, dev)
eng.handle_request({"id": 1, "request": "Create", "learner": {"name": "RegressorPA", "hyperParameters": {"C": 0.01, "trainingMode": "minibatch"}},
                    "preProcessors": [{"name": "PolynomialFeatures", "hyperParameters": {"degree": 2}}, {"name": "StandardScaler"}],
                    "trainingConfiguration": {"protocol": "FGM", "nFeatures": d}})
pool = [R.generate_batch(R.KIND_REGRESSION, d, B, seed=1, batch_index=i).to(dev) for i in range(4)]
for i in range(6):
    eng.train_batch(pool[i % 4], d)
torch.cuda.synchronize()
print("ok")
