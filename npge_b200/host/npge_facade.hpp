// npge_facade.hpp -- the slice of the reference's object model that the
// AnchorFinder processor touches, restated with the standard library only.
//
// The reference's real headers need Boost (program_options, shared_ptr,
// thread), Lua and luabind, none of which exist in this image, so
// gpu_anchor_finder.cpp is compiled and tested against this stand-in.  Names,
// signatures and behaviour follow the reference so that the same .cpp builds
// inside an NPGE checkout with `-DNPGE_B200_REAL_HEADERS`:
//   Sequence      src/model/Sequence.hpp:22-163, Sequence.cpp:151-179 (to_atgcn)
//   Fragment      src/model/Fragment.hpp, Fragment.cpp:27-31,173-183
//   Block         src/model/Block.hpp, Block.cpp:45-50
//   BlockSet      src/model/BlockSet.hpp, BlockSet.cpp:38,62-78
//   Processor     src/algo/Processor.hpp:482,492; Processor.cpp:179-185,652-667
//   Meta          src/algo/Meta.hpp:83-87, Meta.cpp:107-144
// Only what AnchorFinder and its tests use is here; nothing else of NPGE is.
#pragma once

#include <cctype>
#include <cstdint>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace npge {

class Exception : public std::runtime_error {
public:
    explicit Exception(const std::string& what) : std::runtime_error(what) {}
};

// ---- model ---------------------------------------------------------------------
class Sequence {
public:
    virtual ~Sequence() {}
    size_t size() const { return data_.size(); }
    char char_at(size_t index) const { return data_.at(index); }
    const std::string& contents() const { return data_; }
    const std::string& name() const { return name_; }
    void set_name(const std::string& name) { name_ = name; }
    // "GENOME&CHROMOSOME&c" -> "GENOME"; anything else -> "" (Sequence.cpp:193-202)
    std::string genome() const {
        size_t a = name_.find('&');
        if (a == std::string::npos) return "";
        size_t b = name_.find('&', a + 1);
        if (b == std::string::npos || name_.find('&', b + 1) != std::string::npos) return "";
        std::string tail = name_.substr(b + 1);
        return (tail == "c" || tail == "l") ? name_.substr(0, a) : "";
    }
    // Sequence.cpp:151-179: upper-case, IUPAC ambiguity codes -> N, the rest dropped
    static void to_atgcn(std::string& data) {
        size_t d = 0;
        for (size_t f = 0; f < data.size(); f++) {
            char c = char(std::toupper((unsigned char)data[f]));
            if (c == 'A' || c == 'T' || c == 'G' || c == 'C' || c == 'N') {
                data[d++] = c;
            } else if (c != 0 && std::strchr("RYMKWSBVHD", c)) {
                data[d++] = 'N';
            }
        }
        data.resize(d);
    }

protected:
    std::string data_, name_;
};

typedef std::shared_ptr<Sequence> SequencePtr;

class InMemorySequence : public Sequence {
public:
    explicit InMemorySequence(const std::string& raw, const std::string& name = "") {
        data_ = raw;
        to_atgcn(data_);
        name_ = name;
    }
};

inline char complement(char c) {
    switch (c) {
    case 'A': return 'T';
    case 'T': return 'A';
    case 'G': return 'C';
    case 'C': return 'G';
    default: return c;
    }
}

class Fragment {
public:
    Fragment(Sequence* seq, size_t min_pos, size_t max_pos, int ori)
        : seq_(seq), min_pos_(min_pos), max_pos_(max_pos), ori_(ori) {}
    Fragment(SequencePtr seq, size_t min_pos, size_t max_pos, int ori)
        : seq_(seq.get()), min_pos_(min_pos), max_pos_(max_pos), ori_(ori) {}
    Sequence* seq() const { return seq_; }
    size_t min_pos() const { return min_pos_; }
    size_t max_pos() const { return max_pos_; }
    int ori() const { return ori_; }
    size_t length() const { return max_pos_ - min_pos_ + 1; }
    size_t begin_pos() const { return ori_ == 1 ? min_pos_ : max_pos_; }
    // letters in the fragment's own direction (complemented when ori == -1)
    std::string str() const {
        std::string out;
        out.reserve(length());
        for (size_t i = 0; i < length(); i++) {
            out += ori_ == 1 ? seq_->char_at(min_pos_ + i)
                             : complement(seq_->char_at(max_pos_ - i));
        }
        return out;
    }
    // NAME_BEGIN_LAST, last = end position in the fragment's direction (Fragment.cpp:173-183)
    std::string id() const {
        std::ostringstream s;
        long long last = ori_ == 1 ? (long long)max_pos_ : (long long)min_pos_;
        if (min_pos_ == max_pos_ && ori_ == -1) last = -1;  // one-letter reverse fragment
        s << seq_->name() << "_" << begin_pos() << "_" << last;
        return s.str();
    }

private:
    Sequence* seq_;
    size_t min_pos_, max_pos_;
    int ori_;
};

class Block {
public:
    Block() {}
    ~Block() {
        for (Fragment* f : fragments_) delete f;
    }
    Block(const Block&) = delete;
    Block& operator=(const Block&) = delete;
    void insert(Fragment* f) { fragments_.push_back(f); }  // the block owns f
    size_t size() const { return fragments_.size(); }
    bool empty() const { return fragments_.empty(); }
    Fragment* front() const { return fragments_.empty() ? nullptr : fragments_.front(); }
    typedef std::vector<Fragment*>::const_iterator const_iterator;
    const_iterator begin() const { return fragments_.begin(); }
    const_iterator end() const { return fragments_.end(); }

private:
    std::vector<Fragment*> fragments_;
};

class BlockSet {
public:
    BlockSet() {}
    ~BlockSet() { clear_blocks(); }
    BlockSet(const BlockSet&) = delete;
    BlockSet& operator=(const BlockSet&) = delete;
    void add_sequence(SequencePtr seq) { seqs_.insert(seq); }
    // a copy, ordered by pointer like the reference's std::set (BlockSet.cpp:62-64)
    std::vector<SequencePtr> seqs() const {
        return std::vector<SequencePtr>(seqs_.begin(), seqs_.end());
    }
    void insert(Block* block) { blocks_.insert(block); }  // the set owns block
    void detach(Block* block) { blocks_.erase(block); }   // the caller owns it now (BlockSet.cpp)
    bool has(Block* block) const { return blocks_.count(block) != 0; }
    size_t size() const { return blocks_.size(); }
    bool empty() const { return blocks_.empty(); }
    Block* front() const { return blocks_.empty() ? nullptr : *blocks_.begin(); }
    typedef std::set<Block*>::const_iterator const_iterator;
    const_iterator begin() const { return blocks_.begin(); }
    const_iterator end() const { return blocks_.end(); }
    void clear_blocks() {
        for (Block* b : blocks_) delete b;
        blocks_.clear();
    }

private:
    std::set<SequencePtr> seqs_;
    std::set<Block*> blocks_;
};

typedef std::shared_ptr<BlockSet> BlockSetPtr;

inline BlockSetPtr new_bs() { return std::make_shared<BlockSet>(); }

// ---- options: the subset of AnyAs / Processor::add_opt the path uses ------------------
class AnyAs {
public:
    enum Type { NONE, INT, BOOL, DECIMAL, STRING };
    AnyAs() : type_(NONE), i_(0), d_(0) {}
    AnyAs(int v) : type_(INT), i_(v), d_(0) {}
    AnyAs(long long v) : type_(INT), i_(v), d_(0) {}
    AnyAs(bool v) : type_(BOOL), i_(v ? 1 : 0), d_(0) {}
    AnyAs(double v) : type_(DECIMAL), i_(0), d_(v) {}
    AnyAs(const char* v) : type_(STRING), i_(0), d_(0), s_(v) {}
    AnyAs(const std::string& v) : type_(STRING), i_(0), d_(0), s_(v) {}
    Type type() const { return type_; }
    template <typename T> T as() const;
    // text form used on the command line ("--anchor-size 20")
    void from_string(const std::string& text) {
        std::istringstream in(text);
        switch (type_) {
        case INT: in >> i_; break;
        case BOOL: i_ = (text == "1" || text == "true" || text == "yes") ? 1 : 0; return;
        case DECIMAL: in >> d_; break;
        default: s_ = text; return;
        }
        if (in.fail()) throw Exception("bad option value: " + text);
    }

private:
    Type type_;
    long long i_;
    double d_;
    std::string s_;
};

template <> inline int AnyAs::as<int>() const { return int(i_); }
template <> inline long long AnyAs::as<long long>() const { return i_; }
template <> inline bool AnyAs::as<bool>() const { return i_ != 0; }
template <> inline double AnyAs::as<double>() const { return type_ == INT ? double(i_) : d_; }
template <> inline std::string AnyAs::as<std::string>() const { return s_; }

class Processor {
public:
    Processor() : workers_(1), interrupted_(false) {
        // Processor.cpp:179-185: every processor has these
        add_opt("workers", "Number of threads (-1 = number of cores)", 1);
        add_opt("timing", "Measure time for run()", false);
    }
    virtual ~Processor() {}

    void set_block_set(BlockSetPtr bs) { target_ = bs; }
    BlockSetPtr block_set() const { return target_; }
    // the second blockset of two-blockset processors (Processor::other(), "other")
    void set_other(BlockSetPtr bs) { other_ = bs; }
    BlockSetPtr other() const { return other_; }

    bool has_opt(const std::string& name) const { return opts_.count(name) != 0; }
    AnyAs opt_value(const std::string& name) const {
        std::map<std::string, Opt>::const_iterator it = opts_.find(name);
        if (it == opts_.end()) throw Exception("No option with name '" + name + "'");
        return it->second.value;
    }
    void set_opt_value(const std::string& name, const AnyAs& value) {
        std::map<std::string, Opt>::iterator it = opts_.find(name);
        if (it == opts_.end()) throw Exception("No option with name '" + name + "'");
        if (value.type() != it->second.value.type() &&
            !(value.type() == AnyAs::INT && it->second.value.type() == AnyAs::DECIMAL)) {
            throw Exception("Type of value of option '" + name + "' differs");
        }
        it->second.value = value;
        if (name == "workers") workers_ = value.as<int>();
    }
    void set_options(const std::string& text) {  // "--anchor-size 20 --anchor-similar 0"
        std::istringstream in(text);
        std::string key, value;
        while (in >> key) {
            if (key.compare(0, 2, "--") != 0 || !(in >> value)) {
                throw Exception("bad options string: " + text);
            }
            std::string name = key.substr(2);
            AnyAs v = opt_value(name);
            v.from_string(value);
            set_opt_value(name, v);
        }
    }
    std::vector<std::string> opts() const {
        std::vector<std::string> names;
        for (std::map<std::string, Opt>::const_iterator it = opts_.begin();
                it != opts_.end(); ++it) {
            names.push_back(it->first);
        }
        return names;
    }
    const std::string& opt_description(const std::string& name) const {
        return opts_.at(name).description;
    }

    int workers() const { return workers_; }
    void set_workers(int workers) { set_opt_value("workers", workers); }

    std::string key() const { return key_.empty() ? "Processor" : key_; }
    void set_key(const std::string& key) { key_ = key; }
    std::string name() const { return name_impl(); }

    // Processor.cpp:652-667: option rules first, then run_impl unless workers == 0
    void run() const {
        std::string errors;
        for (size_t i = 0; i < rules_.size(); i++) {
            std::string message;
            if (!rules_[i](message)) errors += (errors.empty() ? "" : "; ") + message;
        }
        if (!errors.empty()) {
            throw Exception("Errors in " + key() + "'s options: " + errors);
        }
        if (workers_ != 0 && !interrupted_) run_impl();
    }

protected:
    typedef std::function<bool(std::string&)> Rule;
    void add_opt(const std::string& name, const std::string& description,
                 const AnyAs& default_value) {
        opts_[name] = Opt{description, default_value};
    }
    // global options read their default from the configuration (ANCHOR_SIZE ...);
    // the stand-in has no config file, the defaults are the reference's (src/global/opts.lua)
    void add_gopt(const std::string& name, const std::string& description,
                  const std::string& /*global_name*/, const AnyAs& default_value) {
        add_opt(name, description, default_value);
    }
    void add_opt_rule(const Rule& rule) { rules_.push_back(rule); }
    // ---- the reference's own spellings (src/algo/Processor.hpp:225-238,359-372), so that
    // the NPGE_B200_REAL_HEADERS branch of the processors compiles and runs here too ----
    typedef std::function<bool(std::string&)> OptionsChecker;
    void add_opt_check(const OptionsChecker& checker) { rules_.push_back(checker); }
    // "name op integer", op one of > >= < <= == != (the form every rule on this path has)
    void add_opt_rule(const std::string& rule) {
        std::istringstream in(rule);
        std::string name, op;
        long long rhs = 0;
        if (!(in >> name >> op >> rhs)) throw Exception("bad option rule: " + rule);
        rules_.push_back([this, name, op, rhs, rule](std::string& message) {
            const long long v = opt_value(name).as<int>();
            message = rule;
            if (op == ">") return v > rhs;
            if (op == ">=") return v >= rhs;
            if (op == "<") return v < rhs;
            if (op == "<=") return v <= rhs;
            if (op == "==") return v == rhs;
            if (op == "!=") return v != rhs;
            throw Exception("bad option rule: " + rule);
        });
    }
    // the default comes from the configuration (src/global/opts.lua; CMakeLists.txt:41-46)
    void add_gopt(const std::string& name, const std::string& description,
                  const std::string& global_name) {
        if (global_name == "ANCHOR_SIZE") add_opt(name, description, 20);
        else if (global_name == "ANCHOR_FP") add_opt(name, description, 0.1);
        else if (global_name == "MAX_ANCHOR_FRAGMENTS") add_opt(name, description, 100000);
        else if (global_name == "MAX_MATCHES") add_opt(name, description, 1000);
        else throw Exception("unknown global option " + global_name);
    }
    void declare_bs(const std::string& /*name*/, const std::string& /*description*/) {}
    virtual void run_impl() const = 0;
    virtual const char* name_impl() const { return ""; }

private:
    struct Opt {
        std::string description;
        AnyAs value;
    };
    std::map<std::string, Opt> opts_;
    std::vector<Rule> rules_;
    BlockSetPtr target_, other_;
    std::string key_;
    int workers_;
    bool interrupted_;
};

// ---- Meta: the factory map processors are looked up in (Meta.cpp:107-144) -----------
class Meta {
public:
    typedef std::function<Processor*()> Factory;
    template <typename P>
    void set_processor(const std::string& key, bool overwrite = true) {
        if (!overwrite && factories_.count(key)) return;
        factories_[key] = []() -> Processor* { return new P; };
    }
    bool has(const std::string& key) const { return factories_.count(key) != 0; }
    std::shared_ptr<Processor> get(const std::string& key) const {
        std::map<std::string, Factory>::const_iterator it = factories_.find(key);
        if (it == factories_.end()) throw Exception("No such processor: " + key);
        std::shared_ptr<Processor> p(it->second());
        p->set_key(key);
        return p;
    }

private:
    std::map<std::string, Factory> factories_;
};

}  // namespace npge
