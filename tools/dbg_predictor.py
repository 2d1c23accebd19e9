import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.test_gpu_models import small_model_config, small_weights
from nano_qwen3tts_vllm_b200.llm import PredictorLLM
td, pd, tw, pw = small_weights()
p = PredictorLLM(small_model_config(), state_dict=pw, num_kvcache_blocks=8, max_num_seqs=16, max_model_len=1024,
                 graph_batch_sizes=[1, 2], enforce_eager=("eager" in sys.argv))
r = p.model_runner
r.set_decode_context(2)
out = r.decode_body(2)
torch.cuda.synchronize()
print("ok", out.shape)
