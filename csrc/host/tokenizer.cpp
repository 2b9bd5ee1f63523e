// Native uncased BERT WordPiece tokenizer for batches of flow sentences (host code, no CUDA).
//
// The reference tokenises with HF `DistilBertTokenizer` inside `Dataset.__getitem__`, once per sample per epoch
// (reference client1.py:36-45, 364).  This library tokenises a whole split once, on all cores: clean -> lower-case ->
// split on whitespace and punctuation -> greedy longest-match-first WordPiece -> [CLS] ... [SEP] + [PAD], written
// straight into the int32 [n, max_len] id matrix and the [n] length vector that are uploaded to the GPU.
//
// Scope: the exact BERT algorithm for ASCII text (which is what the flow-sentence template produces).  A sentence
// containing any byte >= 0x80 is not touched; its `fallback` flag is set and the Python implementation
// (data/tokenizer.py, full Unicode: NFD accent stripping, Unicode punctuation / whitespace classes, CJK spacing)
// handles that row, so results are identical to the Python tokenizer for every input.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

struct Tokenizer {
    std::vector<std::string> vocab;                          // owns the strings the map's views point into
    std::unordered_map<std::string_view, int32_t> index;
    int32_t pad_id = 0, unk_id = 100, cls_id = 101, sep_id = 102;
    int lower = 1;
    int max_chars = 100;
};

inline bool is_ascii_punct(unsigned char c) {
    return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126);
}
inline bool is_ws(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
// Unicode category Cc within ASCII: 0x00-0x1F and 0x7F; \t \n \r count as whitespace instead
inline bool is_dropped(unsigned char c) { return c == 0 || ((c < 32 || c == 127) && !is_ws(c)); }

// greedy longest-match-first WordPiece of one word; appends ids, at most `room` of them are kept
inline void wordpiece(const Tokenizer& t, const char* w, int n, std::string& scratch, std::vector<int32_t>& out) {
    if (n > t.max_chars) { out.push_back(t.unk_id); return; }
    const size_t mark = out.size();
    int start = 0;
    while (start < n) {
        int end = n;
        int32_t cur = -1;
        while (start < end) {
            std::string_view piece;
            if (start > 0) {
                scratch.assign("##");
                scratch.append(w + start, end - start);
                piece = scratch;
            } else {
                piece = std::string_view(w, end);
            }
            auto it = t.index.find(piece);
            if (it != t.index.end()) { cur = it->second; break; }
            --end;
        }
        if (cur < 0) {  // no piece matches: the whole word is [UNK]
            out.resize(mark);
            out.push_back(t.unk_id);
            return;
        }
        out.push_back(cur);
        start = end;
    }
}

void encode_one(const Tokenizer& t, const char* s, int64_t len, int max_len, int32_t* ids, int32_t* out_len,
                std::string& word, std::string& scratch, std::vector<int32_t>& body) {
    body.clear();
    const int budget = max_len - 2;
    word.clear();
    auto flush = [&]() {
        if (!word.empty()) {
            if (static_cast<int>(body.size()) < budget) wordpiece(t, word.data(), static_cast<int>(word.size()), scratch, body);
            word.clear();
        }
    };
    for (int64_t i = 0; i < len && static_cast<int>(body.size()) < budget; ++i) {
        const unsigned char c = static_cast<unsigned char>(s[i]);
        if (is_dropped(c)) continue;
        if (is_ws(c)) { flush(); continue; }
        if (is_ascii_punct(c)) {
            flush();
            if (static_cast<int>(body.size()) < budget) {
                const char p = static_cast<char>(c);
                wordpiece(t, &p, 1, scratch, body);
            }
            continue;
        }
        word.push_back(t.lower && c >= 'A' && c <= 'Z' ? static_cast<char>(c + 32) : static_cast<char>(c));
    }
    flush();
    const int nb = std::min<int>(static_cast<int>(body.size()), budget);
    ids[0] = t.cls_id;
    for (int j = 0; j < nb; ++j) ids[1 + j] = body[j];
    ids[1 + nb] = t.sep_id;
    for (int j = nb + 2; j < max_len; ++j) ids[j] = t.pad_id;
    *out_len = nb + 2;
}

}  // namespace

extern "C" {

// vocab_blob: the vocabulary file contents ('\n'-separated entries, entry i has id i)
void* dcad_tok_create(const char* vocab_blob, int64_t blob_len, int lower_case, int max_chars_per_word) {
    auto* t = new Tokenizer();
    t->lower = lower_case;
    t->max_chars = max_chars_per_word > 0 ? max_chars_per_word : 100;
    int64_t start = 0;
    for (int64_t i = 0; i <= blob_len; ++i) {
        if (i == blob_len || vocab_blob[i] == '\n') {
            if (i > start || i < blob_len) t->vocab.emplace_back(vocab_blob + start, static_cast<size_t>(i - start));
            start = i + 1;
        }
    }
    while (!t->vocab.empty() && t->vocab.back().empty()) t->vocab.pop_back();  // trailing newline
    t->index.reserve(t->vocab.size() * 2);
    for (size_t i = 0; i < t->vocab.size(); ++i)  // a repeated entry keeps its LAST id, like the Python dict
        t->index[std::string_view(t->vocab[i])] = static_cast<int32_t>(i);
    auto id = [&](const char* k, int32_t dflt) {
        auto it = t->index.find(std::string_view(k));
        return it == t->index.end() ? dflt : it->second;
    };
    t->pad_id = id("[PAD]", 0); t->unk_id = id("[UNK]", 100); t->cls_id = id("[CLS]", 101); t->sep_id = id("[SEP]", 102);
    return t;
}

void dcad_tok_destroy(void* h) { delete static_cast<Tokenizer*>(h); }

int64_t dcad_tok_vocab_size(void* h) { return static_cast<int64_t>(static_cast<Tokenizer*>(h)->vocab.size()); }

// bytes: all sentences back to back (UTF-8); sentence i = bytes[offsets[i] .. offsets[i+1]).
// ids [n * max_len], lens [n], fallback [n] (1 = contains non-ASCII bytes, row left untouched for the caller).
// Returns the number of rows tokenised natively, or -1 on bad arguments.
int64_t dcad_tok_encode_batch(void* h, const char* bytes, const int64_t* offsets, int64_t n, int max_len, int32_t* ids,
                              int32_t* lens, uint8_t* fallback, int threads) {
    if (!h || n < 0 || max_len < 3) return -1;
    const Tokenizer& t = *static_cast<Tokenizer*>(h);
    if (threads <= 0) threads = static_cast<int>(std::thread::hardware_concurrency());
    threads = std::max(1, std::min<int>(threads, static_cast<int>((n + 255) / 256)));
    std::atomic<int64_t> next{0}, done{0};
    auto work = [&]() {
        std::string word, scratch;
        std::vector<int32_t> body;
        body.reserve(static_cast<size_t>(max_len) + 8);
        int64_t local = 0;
        for (;;) {
            const int64_t b0 = next.fetch_add(256);
            if (b0 >= n) break;
            const int64_t b1 = std::min<int64_t>(n, b0 + 256);
            for (int64_t i = b0; i < b1; ++i) {
                const char* s = bytes + offsets[i];
                const int64_t len = offsets[i + 1] - offsets[i];
                bool ascii = true;
                for (int64_t k = 0; k < len; ++k) if (static_cast<unsigned char>(s[k]) >= 0x80) { ascii = false; break; }
                fallback[i] = ascii ? 0 : 1;
                if (!ascii) continue;
                encode_one(t, s, len, max_len, ids + i * max_len, lens + i, word, scratch, body);
                ++local;
            }
        }
        done.fetch_add(local);
    };
    std::vector<std::thread> pool;
    for (int k = 1; k < threads; ++k) pool.emplace_back(work);
    work();
    for (auto& th : pool) th.join();
    return done.load();
}

}  // extern "C"
