// Test scaffolding for the oracle build only: minimal stand-ins for the protobuf / libvgio
// headers that rpvg's hot-path translation units include transitively. Not product code.
#pragma once
#include <string>
#include <vector>
#include <map>
#include <cstdint>
#include "google/protobuf/util/json_util.h"
namespace vg {
struct AnnValue { std::string s_; double n_=0; const std::string & string_value() const {return s_;} double number_value() const {return n_;} };
struct Annotation { std::map<std::string, AnnValue> f_; const std::map<std::string, AnnValue> & fields() const {return f_;} };
struct Position { int64_t node_id_=0, offset_=0; bool rev_=false; int64_t node_id() const {return node_id_;} int64_t offset() const {return offset_;} bool is_reverse() const {return rev_;} void set_offset(int64_t o){offset_=o;} void set_is_reverse(bool r){rev_=r;} };
struct Edit { int32_t from_=0,to_=0; int32_t from_length() const {return from_;} int32_t to_length() const {return to_;} };
struct Mapping : google::protobuf::Message { Position pos_; std::vector<Edit> edits_; const Position & position() const {return pos_;} Position * mutable_position(){return &pos_;} bool has_position() const {return true;} int edit_size() const {return edits_.size();} const Edit & edit(int i) const {return edits_[i];} Edit * add_edit(){edits_.emplace_back(); return &edits_.back();} };
struct Path : google::protobuf::Message { std::vector<Mapping> m_; int mapping_size() const {return m_.size();} const Mapping & mapping(int i) const {return m_[i];} const std::vector<Mapping> & mapping() const {return m_;} Mapping * add_mapping(){m_.emplace_back(); return &m_.back();} };
struct Alignment : google::protobuf::Message { std::string fld_; Annotation ann_; const std::string & fragment_length_distribution() const {return fld_;} bool has_annotation() const {return !ann_.f_.empty();} const Annotation & annotation() const {return ann_;} std::string seq_, qual_; int32_t score_=0, mq_=0; Path path_; const std::string & sequence() const {return seq_;} const std::string & quality() const {return qual_;} void set_sequence(const std::string&s){seq_=s;} void set_quality(const std::string&s){qual_=s;} int32_t score() const {return score_;} void set_score(int32_t s){score_=s;} int32_t mapping_quality() const {return mq_;} void set_mapping_quality(int32_t s){mq_=s;} const Path & path() const {return path_;} Path * mutable_path(){return &path_;} };
struct Connection { int32_t next_=0, score_=0; int32_t next() const {return next_;} int32_t score() const {return score_;} void set_next(int32_t n){next_=n;} void set_score(int32_t s){score_=s;} };
struct Subpath { Path path_; int32_t score_=0; std::vector<uint32_t> next_; std::vector<Connection> conn_; const Path & path() const {return path_;} Path * mutable_path(){return &path_;} int32_t score() const {return score_;} void set_score(int32_t s){score_=s;} int next_size() const {return next_.size();} uint32_t next(int i) const {return next_[i];} void add_next(uint32_t n){next_.push_back(n);} int connection_size() const {return conn_.size();} const std::vector<Connection> & connection() const {return conn_;} Connection * add_connection(){conn_.emplace_back(); return &conn_.back();} };
struct MultipathAlignment : google::protobuf::Message { Annotation ann_; bool has_annotation() const {return !ann_.f_.empty();} const Annotation & annotation() const {return ann_;} std::string seq_, qual_; int32_t mq_=0; std::vector<Subpath> sp_; std::vector<uint32_t> start_; const std::string & sequence() const {return seq_;} const std::string & quality() const {return qual_;} void set_sequence(const std::string&s){seq_=s;} void set_quality(const std::string&s){qual_=s;} int32_t mapping_quality() const {return mq_;} void set_mapping_quality(int32_t s){mq_=s;} int subpath_size() const {return sp_.size();} const Subpath & subpath(int i) const {return sp_[i];} Subpath * add_subpath(){sp_.emplace_back(); return &sp_.back();} Subpath * mutable_subpath(int i){return &sp_[i];} int start_size() const {return start_.size();} void add_start(uint32_t s){start_.push_back(s);} };
struct Graph : google::protobuf::Message {};
}
