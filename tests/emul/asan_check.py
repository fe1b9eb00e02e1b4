"""TEST INFRASTRUCTURE: the kernels compiled against the fibre-based CUDA stand-in AND AddressSanitizer
(make -C tests/emul asan), driven through the same parity cases: device buffers are plain heap blocks there,
so an out-of-bounds index in a kernel is an ASan report (compute-sanitizer is closed on the GPU pool).
  make -C tests/emul asan && LD_PRELOAD=$(/usr/bin/g++ -print-file-name=libasan.so) \
      ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0 python tests/emul/asan_check.py
Last run (end of round 2): every case below passes with no report."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import parity_cases as pc
L = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'libgpqhe_b200_emul_asan.so')
pc.check_golden(L,'pow2_n64_q438'); print('golden ok', flush=True)
pc.check_he_vs_oracle(L,5,1<<438,4,1<<50,seed=5,levels=[8,1]); print('he ok', flush=True)
pc.check_lazy_ntt_vs_exact(L,13,1<<120); print('lazy13 ok', flush=True)
pc.check_limb_layer_vs_oracle(L,9,1<<100,seed=9); print('limb ok', flush=True)
pc.check_rlsin(L,4,1<<438,1<<50,1,logqub=438); print('rlsin ok', flush=True)
pc.check_gemv(L,5,1<<120,1<<40,2,2,logqub=120); print('gemv ok', flush=True)
pc.check_crt_edges(L); print('edges ok', flush=True)
pc.check_coeff2slot(L,5,1<<120,1<<40,2,2,logqub=120,lanes=3); print('coeff2slot ok', flush=True)
pc.check_he_vs_oracle(L,6,(1<<119)+12345,4,(1<<40)+7,seed=6); print('general modulus ok', flush=True)
pc.check_lazy_ntt_vs_exact(L,16,1<<61); print('lazy16 ok', flush=True)
# round 2: tensor-core CRT stages at 8 / 2 result words and ragged limb counts, fused loaders, fused rescale
pc.check_he_vs_oracle(L,5,1<<500,4,1<<50,seed=3,levels=[9,5,1],logqub=500); print('tc W=8 ok', flush=True)
pc.check_he_vs_oracle(L,5,1<<65,4,1<<20,seed=4,logqub=65); print('tc W=2 ok', flush=True)
pc.check_crt_edges(L,5,1<<500,1<<50,500); print('tc edges ok', flush=True)
pc.check_fused_rescale_ties(L); print('ties ok', flush=True)
pc.check_fused_vs_separate(L,9); print('fused vs separate ok', flush=True)
