"""Constants of the megakernel build (the reference keeps per-task tile configs in core/config.py and tasks/*.py; here
they are compile-time constants of csrc/mega/mega.h, restated for the Python side and the docs)."""
THREADS_PER_CTA = 512          # 16 warps, one CTA per SM (cooperative launch)
MAX_AR_TILES = 32              # tiles of one cross-rank all-reduce op
NUM_KV_SPLITS_DEFAULT = 16     # split-KV factor of the decode attention task
STAGE_ROWS = 24                # weight-stream stage: up to 24 rows of W ...
STAGE_BYTES = STAGE_ROWS * 256 * 2   # ... x 256..768 k-elements = 12 KB, one 3-D TMA box
MAX_STAGE_BUFFERS = 16         # one stage buffer per warp
L2_PREFETCH_ROUNDS = 1         # L2 prefetch distance behind every stage load (buffer rounds)
INT_PER_TASK = 8               # Task record: op, type, tile, dep0, dep0_cnt, dep1, dep1_cnt, out_slot
TRACE_WORDS_PER_TASK = 6       # begin, end, deps satisfied, activations staged, stream consumed, spare
SMEM_LIMIT = 232448            # 227 KB opt-in dynamic shared memory per CTA on sm_100
