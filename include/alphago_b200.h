/*
 * alphago_b200.h -- C ABI of libalphago_b200.so (hand-written sm_100a CUDA, one GPU per process).
 *
 * This is the drop-in boundary for the reference's supervised-learning hot path.  The reference
 * (linghuaj/AlphaGo, a pure-Python project) has no FFI of its own: its boundary is the Python API
 * Preprocess / CNNPolicy / run_training.  The alphago_b200 Python package keeps that API and binds the entry
 * points below with ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 * Each declaration cites the reference interface it replaces (paths relative to the reference).
 *
 * Conventions
 *   - every function returns 0 on success or a negative AG_E* code; ag_last_error() gives the text
 *   - "d_" pointers are DEVICE pointers owned by the caller, "h_" pointers are host pointers
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); no entry point
 *     synchronises the host unless it says so
 *   - points are (x,y), flat index x*size+y; planes are [f][x][y] (AlphaGo/util.py:6-8)
 *   - colours: +1 black, -1 white, 0 empty (AlphaGo/go.py:3-5)
 *   - thread-compatible, not thread-safe: one caller thread per process/GPU
 */
#ifndef ALPHAGO_B200_H
#define ALPHAGO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AG_OK 0
#define AG_EINVAL (-1) /* bad argument */
#define AG_ECUDA (-2)  /* CUDA runtime / driver error */
#define AG_ENOMEM (-3)
#define AG_ESTATE (-4) /* illegal move, wrong object state */

/* feature ids: order of DEFAULT_FEATURES (AlphaGo/preprocessing/preprocessing.py:211-214);
 * AG_F_COLOR is the value net's 49th "side to move is black" plane (AlphaGo/models/value.py:16). */
enum {
    AG_F_BOARD = 0, AG_F_ONES, AG_F_TURNS_SINCE, AG_F_LIBERTIES, AG_F_CAPTURE_SIZE, AG_F_SELF_ATARI_SIZE,
    AG_F_LIBERTIES_AFTER, AG_F_LADDER_CAPTURE, AG_F_LADDER_ESCAPE, AG_F_SENSIBLENESS, AG_F_ZEROS, AG_F_COLOR,
    AG_F_COUNT
};

/* ------------------------------------------------------------------ library */
const char* ag_last_error(void);
int ag_version(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
long long ag_launch_count(void);
/* planes a feature contributes (FEATURES[name]["size"], preprocessing.py:164-209); -1 if unknown */
int ag_feature_planes(int feature);

/* ------------------------------------------------------------------ stage 1: feature encoder
 * Replaces Preprocess.state_to_tensor (preprocessing.py:238-245) and the eleven feature functions
 * it calls (preprocessing.py:9-161, on go.py:63-267) for a batch of n positions of one board size.
 *   d_board  int8  [n][size*size]   GameState.board (go.py:14)
 *   d_age    uint8 [n][size*size]   turns_since input: ag_age_from_history() of (history, board)
 *   d_player int8  [n]              GameState.current_player
 *   d_ko     int16 [n]              GameState.ko as flat index, -1 for None (an INPUT: go.py:322-331
 *                                   can leave a stale ko that cannot be re-derived from the board)
 *   h_features / n_features         feature ids in list order (any subset, order, repeats)
 *   d_out    uint8 [n][planes][size*size], one-hot 0/1 (the reference emits the same values as float64)
 * size in {2..19}; total planes <= 64.  Two launches when ladder planes are asked for (the encoder queues the candidate
 * moves, a second kernel reads them out); their work list lives in a per-device work area of the library that is
 * allocated on the first such call and grown when a call brings more positions than any before (cudaMalloc / cudaFree:
 * those calls synchronise the device); all other calls allocate nothing.  At most four encodes in flight per device. */
int ag_encode(const int8_t* d_board, const uint8_t* d_age, const int8_t* d_player, const int16_t* d_ko,
              int n, int size, const int32_t* h_features, int n_features, uint8_t* d_out, void* stream);

/* ------------------------------------------------------------------ host rules engine
 * Backs alphago_b200.go.GameState (the product's mirror of AlphaGo/go.py:9-349) and the synthetic
 * random-playout generator.  Pure host code. */
typedef struct ag_state ag_state;
ag_state* ag_state_new(int size);                                  /* go.py:13-45 */
ag_state* ag_state_copy(const ag_state* s);                        /* go.py:159-179 (history de-aliased) */
void ag_state_free(ag_state* s);
int ag_state_do_move(ag_state* s, int point, int colour);          /* go.py:295-349; point -1 = pass; colour 0 = side to move; AG_ESTATE = IllegalMove */
int ag_state_is_legal(ag_state* s, int point);                     /* go.py:204-216 */
int ag_state_is_eyeish(ag_state* s, int point, int owner);         /* go.py:218-228 */
int ag_state_is_eye(ag_state* s, int point, int owner);            /* go.py:230-259 */
int ag_state_legal_moves(ag_state* s, int include_eyes, int16_t* h_out); /* go.py:261-267; returns count */
int ag_state_winner(const ag_state* s, double komi);               /* go.py:269-293 */
/* scalar fields: 0 size, 1 current_player, 2 ko, 3 len(history), 4 is_end_of_game, 5 turns_played,
 * 6 num_black_prisoners, 7 num_white_prisoners, 8 passes_black, 9 passes_white */
int ag_state_get(const ag_state* s, int field);
int ag_state_set_player(ag_state* s, int colour);
int ag_state_set_ko(ag_state* s, int point);
void ag_state_board(const ag_state* s, int8_t* h_board);
void ag_state_history(const ag_state* s, int16_t* h_hist);
/* per-point group id (-1 empty), stones and liberties of the point's group (liberty_counts, go.py:40) */
void ag_state_groups(ag_state* s, int16_t* h_gid, int16_t* h_stones, int16_t* h_libs);
/* encoder inputs of this state */
void ag_state_pack(const ag_state* s, int8_t* h_board, uint8_t* h_age, int8_t* h_player, int16_t* h_ko);
/* A whole game record in one call: for i in [0, n_moves) -- if h_points[i] is not a pass (-1), pack the position as
 * ag_state_pack does into slot k of the outputs (k counts up from 0), then play h_points[i] for h_colours[i] (+1 black,
 * -1 white).  That is the loop of game_converter.convert_game over util.sgf_iter_states (game_converter.py:35-40,
 * util.py:64-83) without a Python round trip per move.  Outputs hold up to n_moves positions of size*size points;
 * *n_packed = positions written.  AG_ESTATE on an illegal move: the pairs up to and including that move's are kept,
 * as the reference keeps what it had written (game_converter.py:111-112). */
int ag_state_replay_pack(ag_state* s, int n_moves, const int16_t* h_points, const int8_t* h_colours, int8_t* h_boards,
                         uint8_t* h_ages, int8_t* h_players, int16_t* h_kos, int* n_packed);
/* the literal turns_since walk (preprocessing.py:20-43) for any (history, board) pair */
void ag_age_from_history(int size, const int8_t* h_board, const int16_t* h_hist, int n_hist, uint8_t* h_age);
/* n positions from seeded random playouts, uniform over get_legal_moves(include_eyes=False)
 * (AlphaGo/ai.py:15), `per_game` snapshots per game at move numbers uniform in [0, max_moves] */
int ag_playout_positions(uint64_t seed, int n, int size, int max_moves, int per_game,
                         int8_t* h_board, uint8_t* h_age, int8_t* h_player, int16_t* h_ko);

/* ------------------------------------------------------------------ stage 2: conv-stack network
 * One object per GPU holds the weights, activations and gradients of a CNNPolicy / CNNValue net
 * (AlphaGo/models/policy.py:93-156, AlphaGo/models/value.py:12-28): n_conv KxK 'same' ReLU convs
 * with a common filter count (widths odd, <= 5), then a 1x1 conv to one plane and
 *   head_kind 0: Flatten + softmax over board*board moves (+ optional per-position bias)
 *   head_kind 1: Flatten + Dense(256, linear) + Dense(1, tanh)
 * bf16 tensor-core math with fp32 accumulation, fp32 master weights.  All device memory is
 * allocated here (sized for max_batch); later calls allocate nothing.  NULL + ag_last_error() on failure. */
typedef struct ag_net ag_net;
ag_net* ag_net_create(int board, int in_planes, int n_conv, const int* h_filters, const int* h_widths,
                      int head_kind, int max_batch);                /* policy.py:93-156 create_network */
void ag_net_destroy(ag_net* net);
/* 0 cin, 1 cout, 2 width, 3 padded cin, 4 padded cout; how the layer is launched (no counterpart in the reference):
 * 5 = its forward runs inside the one chained launch, 6 = its data gradient does, 7 = its weight gradient does (known
 * after a forward_loss), 8 = chained launches the device refused so far (whole net; each was answered by the per-layer
 * kernels), 9 = split-K factor of its weight gradient */
int ag_net_layer_info(const ag_net* net, int layer, int field);
/* conv weights, cross-correlation filters in Keras 'th' order w[co][ci][kx][ky], b[co] (host, synchronous);
 * Keras model.get_weights / set_weights (tests/test_policy.py:42-51, policy.py:165-167,188-190) */
int ag_net_set_conv(ag_net* net, int layer, const float* h_w, const float* h_b);
int ag_net_get_conv(ag_net* net, int layer, float* h_w, float* h_b);
/* head parameters: which 0 = 1x1 conv (w[C], b[1]); 1 = position bias (w[board*board]; setting it enables it);
 * 2 = value Dense(256) (w[256][board*board], b[256]); 3 = value Dense(1) (w[256], b[1]) */
int ag_net_set_head(ag_net* net, int which, const float* h_w, const float* h_b);
int ag_net_get_head(ag_net* net, int which, float* h_w, float* h_b);
/* CNNPolicy.forward (policy.py:26-42): d_planes uint8 [n][in_planes][board][board] (the encoder's output /
 * the HDF5 `states` layout) -> d_out fp32 [n][board*board] probabilities (head 0) or [n] values (head 1) */
int ag_net_forward(ag_net* net, const uint8_t* d_planes, int n, float* d_out, void* stream);
/* forward + categorical cross-entropy + accuracy + head backward (supervised_policy_trainer.py:200).
 * d_labels int32 [n] = x*board+y.  d_loss_acc fp32[2] is ACCUMULATED: sum of losses, number correct.
 * d_probs may be NULL.  Gradients are means over n. */
int ag_net_forward_loss(ag_net* net, const uint8_t* d_planes, const int32_t* d_labels, int n, float* d_probs,
                        float* d_loss_acc, void* stream);
/* backward of conv layer `layer`; call for n_conv-1 ... 0 after ag_net_forward_loss (fit_generator's backprop,
 * supervised_policy_trainer.py:207-213).  The layer's slice of the gradient buffer is final when it returns. */
int ag_net_backward_layer(ag_net* net, int layer, int n, void* stream);
/* the same for layers hi .. lo in one call (descending, non-overlapping ranges starting at n_conv-1): lets runs of equal
 * layers share one chained launch; what data-parallel training calls per gradient bucket.  Replaces the same Keras
 * backward pass as ag_net_backward_layer (supervised_policy_trainer.py:207-213). */
int ag_net_backward_range(ag_net* net, int hi, int lo, int n, void* stream);
/* SMs the weight-gradient launches may fill (0 = all).  No counterpart in the reference (Keras/Theano schedule their own
 * kernels); data-parallel training leaves a few SMs to the gradient all-reduce that runs beside the backward pass. */
int ag_net_set_sm_budget(ag_net* net, int sms);
/* Overlap of the gradient exchange with the backward pass (no counterpart in the reference, which trains on one device).
 * After ag_net_backward_range: *d_counter is a device word that reaches *target (wrap-around >=) once the weight and bias
 * gradients of `layer` are final -- inside the chained launch, while the layers below are still being computed.
 * AG_ESTATE if the layer does not run chained (then its gradient is final when the stream reaches the end of the call).
 * ag_stream_wait_value() makes a stream wait for such a word (cuStreamWaitValue32): what SLTrainer puts in front of each
 * bucket's all-reduce. */
int ag_net_wgrad_done(ag_net* net, int layer, unsigned int** d_counter, unsigned int* target);
int ag_stream_wait_value(void* stream, const unsigned int* d_addr, unsigned int value);
/* ag_net_backward_range over all layers queues the first layer's weight gradient (its own kernels) AHEAD of the chained
 * launch of the others; this makes `stream` wait for exactly that point, so the first layer's gradient can be exchanged
 * while the chain runs.  AG_ESTATE when the latest backward call did not do so. */
int ag_net_wait_first_layer(ag_net* net, void* stream);
/* Gradient exchange over NVLink peer memory (no counterpart in the reference): a two-shot sum all-reduce of ranges of the
 * buffer ag_net_grads() returns, for the ranks of ONE node, one process per GPU.  ag_xchg_create() makes this rank's end and
 * fills h_handle (144 bytes: CUDA IPC handles of the buffer's allocation and of a signal block, the buffer's offset);
 * the caller all-gathers the records of all ranks (any transport) and passes them, in rank order, to ag_xchg_connect().
 * ag_xchg_allreduce(start, count) then sums buf[start : start+count) over the ranks in place -- every rank must make the
 * same calls in the same order; `stream` order (or a gate such as ag_stream_wait_value) must make this rank's part of the
 * range final before the kernel starts.  The owner of a slice reduces it in rank order and everybody copies its result:
 * all ranks end with bit-identical sums.  `ctas` CTAs (32 is plenty) of 256 threads and no shared memory: they fit next
 * to the network's kernels on the same SMs.  Peer memory is read with ordinary loads only. */
typedef struct ag_xchg ag_xchg;
ag_xchg* ag_xchg_create(int rank, int world, float* d_buf, long long count, int ctas, unsigned char* h_handle);
int ag_xchg_connect(ag_xchg* x, const unsigned char* h_all);
int ag_xchg_allreduce(ag_xchg* x, long long start, long long count, void* stream);
/* makes `stream` wait until all ranks have finished reading this rank's buffer in the exchanges issued so far; call it on
 * the stream that overwrites the buffer next (the following backward pass), before it does */
int ag_xchg_wait_readers(ag_xchg* x, void* stream);
void ag_xchg_destroy(ag_xchg* x);
/* loss differentiated by ag_net_forward_loss: 0 = categorical cross-entropy (compile(loss='categorical_crossentropy'),
 * supervised_policy_trainer.py:200), 1 = Keras binary_crossentropy on the softmax outputs
 * (compile(loss='binary_crossentropy'), reinforcement_policy_trainer.py:109) */
int ag_net_set_loss(ag_net* net, int kind);
/* the flat fp32 gradient buffer (device) data-parallel training all-reduces; h_offsets gets n_conv+2 offsets */
int ag_net_grads(ag_net* net, float** d_grads, long long* count, long long* h_offsets);
/* fused SGD(+momentum) over all parameters; gradients are multiplied by grad_scale first
 * (keras.optimizers.SGD as used at supervised_policy_trainer.py:199; lr already decayed by the caller) */
int ag_net_sgd(ag_net* net, float lr, float momentum, float grad_scale, void* stream);
/* single-GPU step: forward_loss + every backward_layer + sgd */
int ag_net_train_step(ag_net* net, const uint8_t* d_planes, const int32_t* d_labels, int n, float lr, float momentum,
                      float* d_loss_acc, void* stream);
/* Stand-alone weight-gradient op on caller-owned buffers (the backward of one Keras Convolution2D layer,
 * policy.py:119-142, as run inside fit_generator, supervised_policy_trainer.py:207-213):
 *   d_dw[tap][co][ci] = sum_m d_g[m - shift(tap)][co] * d_x[m][ci],  shift(kx,ky) = (kx-w/2)*row_stride + (ky-w/2)
 * d_g bf16 [rows][ldout] = gradient w.r.t. the layer's pre-activation output, d_x bf16 [rows][kin] = its input,
 * both "padded pixel rows" (>= w/2 zero slots after every board row and zero rows after
 * every image); ldout, kin multiples of 64 (kin <= 256); width odd <= 5; d_dw fp32 [width*width][ldout][kin].
 * d_scratch holds the split-K partials: ag_wgrad_scratch_bytes() is the preferred size, anything down to one
 * partial (ag_wgrad_partial_bytes(): whole accumulator tiles, a little over width*width*ldout*kin*4 bytes) works
 * with fewer splits.  Deterministic (fixed-order reduction). */
size_t ag_wgrad_scratch_bytes(int width, int ldout, int kin, long long rows);
size_t ag_wgrad_partial_bytes(int width, int ldout, int kin);
int ag_wgrad(const void* d_g, const void* d_x, long long rows, int ldout, int kin, int width, int row_stride,
             float* d_scratch, size_t scratch_bytes, float* d_dw, void* stream);
/* ag_wgrad for callers that must not allocate.  (No counterpart in the reference: nothing there manages device
 * memory.)  Same arguments and result as ag_wgrad without the scratch pair: the split-K partials go to a fixed
 * array built into the library image, ag_wgrad_static_max_bytes() = 16 MiB, which every GPU that loads the
 * library holds until the process exits.  Meant for code that runs inside someone else's memory plan -- a
 * pre-sized pool that must stay untouched, a serving loop where cudaMalloc would stall or fail, a process with a
 * fixed device-memory budget.  It calls no cudaMalloc/cudaFree and no allocating library routine.
 *   - a problem whose single partial (ag_wgrad_partial_bytes()) is larger than the array is refused with
 *     AG_EINVAL and a message naming both sizes; smaller problems use as many splits as fit
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
 *     current); make that first call before starting any stream capture
 *   - all calls share the one array and it keeps the previous call's partials; the kernel overwrites every
 *     element it later reads, so nothing needs clearing -- but two calls whose kernels overlap on the GPU
 *     (different streams, or different host threads) race on it: keep them ordered on one stream */
size_t ag_wgrad_static_max_bytes(void);
int ag_wgrad_static(const void* d_g, const void* d_x, long long rows, int ldout, int kin, int width, int row_stride,
                    float* d_dw, void* stream);
/* Device-resident self-play (the mini-batch of games of reinforcement_policy_trainer.py:16-76 and the move choice of
 * ProbabilisticPolicyPlayer, ai.py:38-68, without a host round trip per ply).  `games` positions of one board size live
 * in the encoder's input arrays (d_board / d_age / d_player / d_ko: feed them to ag_encode as they are) plus
 *   d_flags  uint8 [games]     bit 0 = is_end_of_game (go.py:345-348), bit 1 = the previous move was a pass
 *   d_passes int32 [games][2]  passes_black, passes_white (go.py:341-344)
 * One call plays one ply in every game that is not over: it samples a move with probability proportional to
 * d_probs[g][p] ^ beta over the points where d_mask[g * mask_stride + p] != 0 (the encoder's `sensibleness` plane of that
 * position: legal and not the mover's own eye, ai.py:15,38), using the uniform number d_uniform[g] in [0, 1) -- no such point
 * means pass (ai.py:49) -- and applies it with the reference's rules (go.py:295-349: captures, ko, pass bookkeeping).
 * d_moves int16 [games]: the move played (x*size+y), -1 for a pass, -2 if the game was already over. */
int ag_selfplay_step(int8_t* d_board, uint8_t* d_age, int8_t* d_player, int16_t* d_ko, uint8_t* d_flags, int32_t* d_passes,
                     const float* d_probs, const uint8_t* d_mask, long long mask_stride, const float* d_uniform, float beta,
                     int games, int size, int16_t* d_moves, void* stream);
/* GameState.get_winner (go.py:269-293) for every game: +1 black, -1 white, 0 tie */
int ag_selfplay_winner(const int8_t* d_board, const int32_t* d_passes, int games, int size, float komi, int8_t* d_winner, void* stream);
/* LZF byte streams on the host (no GPU work): the chunk filter the reference's converter asks h5py for
 * (compression="lzf", game_converter.py:73-74,81-82).  decompress returns the number of bytes written, or -1 on a
 * corrupt stream / too small an output buffer; compress returns the stream's length, or 0 when it does not fit
 * out_cap (the caller then stores the chunk unfiltered, as the HDF5 filter does). */
long long ag_lzf_decompress(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap);
long long ag_lzf_compress(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_cap);
/* Minibatch assembly (shuffled_hdf5_batch_generator + BOARD_TRANSFORMATIONS + one_hot_action,
 * supervised_policy_trainer.py:10-40,71-80): for b in [0, batch): sample d_index[b] of the dataset
 * (d_states uint8 [N][planes][size][size], d_actions uint8 [N][2] as (x, y)) under dihedral transform
 * d_transform[b] in 0..7 (the order of BOARD_TRANSFORMATIONS) -> d_out uint8 [batch][planes][size][size]
 * and d_labels int32 [batch] = flat index of the transformed action. */
int ag_augment(const uint8_t* d_states, const uint8_t* d_actions, const int64_t* d_index, const uint8_t* d_transform,
               int batch, int planes, int size, uint8_t* d_out, int32_t* d_labels, void* stream);
/* measurement hook for bench.py's per-kernel roofline: re-run one kernel of the step on the buffers in place.
 * mode 0 = forward conv of `layer`, 1 = its data gradient, 2 = its weight gradient, 3 = the whole forward conv stack as
 * one chained launch, 4 = the whole data-gradient chain, 5 = the weight gradients of every hidden layer as one chained
 * launch (3, 4 and 5 ignore `layer`; AG_ESTATE when the net runs unchained) */
int ag_net_run_layer(ag_net* net, int layer, int mode, int n, void* stream);
/* test hook (no counterpart in the reference): `ctas` CTAs that each occupy one SM for `microseconds` (<= 5 s) on `stream`,
 * to put the network's launches next to a kernel that holds SMs */
int ag_debug_hold_sms(int ctas, long long microseconds, void* stream);
/* test/debug readback (synchronous, allocates): kind 0 = activations of conv layer, 1 = its pre-activation
 * gradient, as fp32 [n][cout][board][board] on the host */
int ag_net_debug_read(ag_net* net, int layer, int kind, int n, float* h_out);

#ifdef __cplusplus
}
#endif
#endif /* ALPHAGO_B200_H */
