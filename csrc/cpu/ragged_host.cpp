// Host runtime of the FastGen ragged engine: blocked KV-cache allocator, sequence table and the per-forward batch
// descriptor builder. One call per `put()` turns the scheduled (uid, new tokens) pairs into every integer array the
// device kernels consume (token -> sequence row / position, block table, the decode / prefill / short-chunk partition
// of csrc/ragged_ops.cu and csrc/attention_fwd.cu) inside one caller-provided (pinned) buffer, so the Python side does
// a single host-to-device copy instead of O(tokens) interpreter work.
//
// Reference: inference/v2/ragged/blocked_allocator.py (free list), sequence_descriptor.py, manager.py (DSStateManager),
// ragged_wrapper.py (RaggedBatchWrapper: inflight descriptors + kv pointers in a pinned "fast host buffer",
// ragged/csrc/fast_host_buffer.cu), kernels/ragged_ops/atom_builder (C++ work decomposition).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <unordered_map>
#include <vector>

namespace {

struct Sequence {
    int64_t seen = 0;          // tokens whose KV is already in the cache
    std::vector<int> blocks;   // KV blocks owned, in position order
};

struct RaggedState {
    int num_blocks, block_size;
    std::vector<int> next;     // free list: next[b] = following free block, -1 = allocated
    int head, free_blocks;
    std::unordered_map<int64_t, Sequence> seqs;
    int max_tracked;
    RaggedState(int nb, int bs, int mt) : num_blocks(nb), block_size(bs), next(nb), head(0), free_blocks(nb), max_tracked(mt)
    {
        for (int i = 0; i < nb; i++) next[i] = i + 1;
    }
    bool allocate(int n, std::vector<int>& out)
    {
        if (n > free_blocks) return false;
        for (int i = 0; i < n; i++) {
            const int b = head;
            head = next[b];
            next[b] = -1;
            free_blocks--;
            out.push_back(b);
        }
        return true;
    }
    int release(const int* blocks, int n)
    {
        for (int i = 0; i < n; i++) {
            const int b = blocks[i];
            if (b < 0 || b >= num_blocks) return -1;   // invalid block
            if (next[b] != -1) return -2;              // double free
            next[b] = head;
            head = b;
            free_blocks++;
        }
        return 0;
    }
};

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

}  // namespace

extern "C" {

void* dsb200_ragged_create(int num_blocks, int block_size, int max_tracked_sequences)
{
    if (num_blocks < 1 || block_size < 1) return nullptr;
    return new RaggedState(num_blocks, block_size, max_tracked_sequences);
}

void dsb200_ragged_destroy(void* h) { delete static_cast<RaggedState*>(h); }

int dsb200_ragged_free_blocks(void* h) { return static_cast<RaggedState*>(h)->free_blocks; }
int dsb200_ragged_tracked(void* h) { return (int)static_cast<RaggedState*>(h)->seqs.size(); }

// seen tokens of a sequence, -1 if unknown
int64_t dsb200_ragged_seen(void* h, int64_t uid)
{
    auto* st = static_cast<RaggedState*>(h);
    auto it = st->seqs.find(uid);
    return it == st->seqs.end() ? -1 : it->second.seen;
}

// blocks a sequence would have to get to take `new_tokens` more tokens (does not allocate)
int dsb200_ragged_blocks_needed(void* h, int64_t uid, int new_tokens)
{
    auto* st = static_cast<RaggedState*>(h);
    auto it = st->seqs.find(uid);
    const int64_t seen = it == st->seqs.end() ? 0 : it->second.seen;
    const int have = it == st->seqs.end() ? 0 : (int)it->second.blocks.size();
    return std::max(0, cdiv(seen + new_tokens, st->block_size) - have);
}

// copies the block list of a sequence into out (capacity cap); returns the number of blocks, -1 if unknown
int dsb200_ragged_blocks(void* h, int64_t uid, int* out, int cap)
{
    auto* st = static_cast<RaggedState*>(h);
    auto it = st->seqs.find(uid);
    if (it == st->seqs.end()) return -1;
    const int n = (int)it->second.blocks.size();
    for (int i = 0; i < n && i < cap; i++) out[i] = it->second.blocks[i];
    return n;
}

int dsb200_ragged_flush(void* h, int64_t uid)
{
    auto* st = static_cast<RaggedState*>(h);
    auto it = st->seqs.find(uid);
    if (it == st->seqs.end()) return 0;
    const int rc = st->release(it->second.blocks.data(), (int)it->second.blocks.size());
    st->seqs.erase(it);
    return rc;
}

// Schedules one ragged forward: for every (uid, n_tokens) creates the sequence if new, grows its block list, and writes
// into `buf` (int32, capacity `cap` entries), in this order:
//   tok_seq[T] tok_pos[T] table[n_seq * width]
//   prefill: cu_q[n_p] q_len[n_p] kv_len[n_p] rows[n_p]              (sequences with >= prefill_min new tokens)
//   decode : rows[n_d] kv_len[n_d] seqrow[n_d]                      (sequences with exactly 1 new token)
//   rest   : rows[T_r] seq[T_r] pos[T_r]                            (tokens of all other sequences)
// header[16]: T, n_seq, width, n_p, max_q, n_d, max_ctx_decode, T_r, max_ctx_rest, max_ctx, offsets of the three groups
// (header[10..12]), total ints used (header[13]). `commit` != 0 advances the seen-token counters (post_forward).
// Returns 0; -1 not enough KV blocks (nothing is changed); -2 too many tracked sequences; -3 buffer too small.
int dsb200_ragged_schedule(void* h, int n_seq, const int64_t* uids, const int* n_tokens, int prefill_min, int* buf,
                           int64_t cap, int* header, int commit)
{
    auto* st = static_cast<RaggedState*>(h);
    // admission: total blocks needed
    int need_total = 0, new_seqs = 0;
    for (int s = 0; s < n_seq; s++) {
        need_total += dsb200_ragged_blocks_needed(h, uids[s], n_tokens[s]);
        if (st->seqs.find(uids[s]) == st->seqs.end()) new_seqs++;
    }
    if (need_total > st->free_blocks) return -1;
    if ((int)st->seqs.size() + new_seqs > st->max_tracked) return -2;
    int T = 0, width = 1;
    std::vector<Sequence*> sq(n_seq);
    for (int s = 0; s < n_seq; s++) {
        Sequence& q = st->seqs[uids[s]];
        const int need = std::max(0, cdiv(q.seen + n_tokens[s], st->block_size) - (int)q.blocks.size());
        st->allocate(need, q.blocks);
        sq[s] = &q;
        T += n_tokens[s];
        width = std::max(width, (int)q.blocks.size());
    }
    int n_p = 0, n_d = 0, T_r = 0;
    for (int s = 0; s < n_seq; s++) {
        if (n_tokens[s] >= prefill_min) n_p++;
        else if (n_tokens[s] == 1) n_d++;
        else T_r += n_tokens[s];
    }
    const int64_t total = 2LL * T + (int64_t)n_seq * width + 4LL * n_p + 3LL * n_d + 3LL * T_r;
    if (total > cap) return -3;
    int* tok_seq = buf;
    int* tok_pos = buf + T;
    int* table = buf + 2 * T;
    int* pf = table + (int64_t)n_seq * width;
    int* dc = pf + 4 * n_p;
    int* rs = dc + 3 * n_d;
    std::memset(table, 0, sizeof(int) * (size_t)n_seq * width);
    int t = 0, ip = 0, id = 0, ir = 0, max_q = 0, max_ctx_d = 0, max_ctx_r = 0, max_ctx = 0;
    for (int s = 0; s < n_seq; s++) {
        Sequence& q = *sq[s];
        const int n = n_tokens[s];
        const int seen = (int)q.seen;
        std::copy(q.blocks.begin(), q.blocks.end(), table + (int64_t)s * width);
        for (int i = 0; i < n; i++) {
            tok_seq[t + i] = s;
            tok_pos[t + i] = seen + i;
        }
        max_ctx = std::max(max_ctx, seen + n);
        if (n >= prefill_min) {
            pf[ip] = t;
            pf[n_p + ip] = n;
            pf[2 * n_p + ip] = seen + n;
            pf[3 * n_p + ip] = s;
            max_q = std::max(max_q, n);
            ip++;
        } else if (n == 1) {
            dc[id] = t;
            dc[n_d + id] = seen + 1;
            dc[2 * n_d + id] = s;
            max_ctx_d = std::max(max_ctx_d, seen + 1);
            id++;
        } else {
            for (int i = 0; i < n; i++) {
                rs[ir + i] = t + i;
                rs[T_r + ir + i] = s;
                rs[2 * T_r + ir + i] = seen + i;
            }
            max_ctx_r = std::max(max_ctx_r, seen + n);
            ir += n;
        }
        t += n;
        if (commit) q.seen += n;
    }
    header[0] = T, header[1] = n_seq, header[2] = width, header[3] = n_p, header[4] = max_q, header[5] = n_d;
    header[6] = max_ctx_d, header[7] = T_r, header[8] = max_ctx_r, header[9] = max_ctx;
    header[10] = (int)(pf - buf), header[11] = (int)(dc - buf), header[12] = (int)(rs - buf), header[13] = (int)total;
    return 0;
}

}  // extern "C"
