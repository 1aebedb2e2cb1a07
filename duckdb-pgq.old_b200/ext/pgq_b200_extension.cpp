// pgq_b200.duckdb_extension — DuckDB host shim of the B200 graph hot path.
//
// Registers the DuckPGQ scalar functions (same SQL names, argument vectors and result vectors;
// SURVEY.md §8(a) a2-a7) through DuckDB's scalar-function API and forwards every chunk to the
// C ABI in include/pgq_b200.h.  Nothing here computes: no CPU fallback exists, a missing GPU
// or library surfaces as a duckdb::Exception with the C ABI's message.
//
// Contracts followed (all in /root/reference):
//   plugin ABI  <basename>_init / <basename>_version   src/execution/operator/helper/physical_load.cpp:36-65
//   registration via Catalog::CreateFunction            extension/fts/fts-extension.cpp:51-74
//   scalar_function_t / bind_scalar_function_t          src/include/duckdb/function/scalar_function.hpp:24-27
//   one call = one DataChunk of <= 1024 rows            src/execution/expression_executor/execute_function.cpp:19-40
//   Orrify / VectorData                                 src/include/duckdb/common/types/vector.hpp:22-27,116-119
//   LIST result fill (Reserve / GetEntry / SetListSize) src/function/scalar/list/range.cpp:118-149
//   NULL in / out                                       src/include/duckdb/common/types/validity_mask.hpp:151-156,190-196
//   has_side_effects = true (shared state)              src/function/scalar/sequence/nextval.cpp:137-138
#define DUCKDB_BUILD_LOADABLE_EXTENSION
#include "duckdb.hpp"
#include "duckdb/catalog/catalog.hpp"
#include "duckdb/common/exception.hpp"
#include "duckdb/function/scalar_function.hpp"
#include "duckdb/main/client_context.hpp"
#include "duckdb/parser/parsed_data/create_scalar_function_info.hpp"
#include "duckdb/planner/expression/bound_function_expression.hpp"

#include <cstdlib>
#include <map>
#include <mutex>
#include <vector>

#include "../../include/pgq_b200.h"

namespace duckdb {

// ---- process-wide GPU context + (connection, csr_id) -> device csr id -----------------------------
struct PgqCsrSlot {
	int32_t dev_id = -1;
	bool vertex_started = false; // CsrInitializeVertex done (first create_csr_vertex row seen)
	bool edge_started = false;   // CsrInitializeEdge done (first create_csr_edge row seen)
	bool finalized = false;
	int64_t num_edges = 0, rows_added = 0;
	// 7-argument create_csr_edge: weights of the rows seen so far (raw 8-byte values, input-row order)
	int wtype = 0;
	std::vector<uint64_t> weights;
};

struct PgqState {
	std::mutex mu;
	pgq_ctx_t *ctx = nullptr;
	int32_t next_dev_id = 1;
	std::map<std::pair<ClientContext *, int32_t>, PgqCsrSlot> slots;

	[[noreturn]] static void Fail(const char *what, int rc) {
		const char *msg = pgq_last_error();
		string text = string("pgq_b200: ") + what + ": " + (msg && *msg ? msg : "error") + " (code " +
		              std::to_string(rc) + ")";
		if (rc == PGQ_ERR_OOM) throw OutOfMemoryException(text);
		if (rc == PGQ_ERR_INVALID || rc == PGQ_ERR_MISMATCH) throw InvalidInputException(text);
		throw InternalException(text);
	}
	static void Check(const char *what, int rc) {
		if (rc != PGQ_OK) Fail(what, rc);
	}
	// caller holds mu: K1-K3 run now; weights (if the rows carried any) go into the CSR slots
	void Finish(PgqCsrSlot &s) {
		Check("csr finalize", pgq_csr_finalize(Ctx(), s.dev_id));
		s.finalized = true;
		if (s.wtype) {
			Check("csr weights", pgq_csr_set_weights(Ctx(), s.dev_id, s.wtype, s.weights.data(), 0,
			                                         (int64_t)s.weights.size()));
			s.weights.clear();
			s.weights.shrink_to_fit();
		}
	}
	// caller holds mu
	pgq_ctx_t *Ctx() {
		if (!ctx) {
			const char *dev = std::getenv("PGQ_B200_DEVICE");
			Check("pgq_init", pgq_init(dev ? std::atoi(dev) : 0, &ctx));
		}
		return ctx;
	}
	PgqCsrSlot &Slot(ClientContext *cc, int32_t csr_id) {
		auto key = std::make_pair(cc, csr_id);
		auto it = slots.find(key);
		if (it == slots.end()) {
			PgqCsrSlot s;
			s.dev_id = next_dev_id++;
			it = slots.insert(std::make_pair(key, s)).first;
		}
		return it->second;
	}
	// queries finalize lazily when the edge rows stopped short of an explicit finalize
	PgqCsrSlot &ReadySlot(ClientContext *cc, int32_t csr_id) {
		auto it = slots.find(std::make_pair(cc, csr_id));
		if (it == slots.end() || !it->second.vertex_started) {
			throw InvalidInputException("pgq_b200: csr " + std::to_string(csr_id) +
			                            " does not exist (run create_csr_vertex / create_csr_edge first)");
		}
		PgqCsrSlot &s = it->second;
		if (!s.finalized) {
			Finish(s);
		}
		return s;
	}
};

static PgqState &State() {
	static PgqState st;
	return st;
}

// bind data: the connection the CSR map is keyed by (cf. NextvalBindData capturing the context,
// src/function/scalar/sequence/nextval.cpp:15-27)
struct PgqBindData : public FunctionData {
	explicit PgqBindData(ClientContext &context) : context(context) {
	}
	ClientContext &context;
	unique_ptr<FunctionData> Copy() override {
		return make_unique<PgqBindData>(context);
	}
};

static unique_ptr<FunctionData> PgqBind(ClientContext &context, ScalarFunction &, vector<unique_ptr<Expression>> &) {
	return make_unique<PgqBindData>(context);
}

static ClientContext *ContextOf(ExpressionState &state) {
	auto &expr = (BoundFunctionExpression &)state.expr;
	return &((PgqBindData &)*expr.bind_info).context;
}

// ---- argument helpers --------------------------------------------------------------------------------
// constant-per-call arguments (csr_id, v_size, num_edges) are read from row 0 like the reference does
static int64_t FirstBigint(DataChunk &args, idx_t col) {
	VectorData vd;
	args.data[col].Orrify(args.size(), vd);
	auto idx = vd.sel->get_index(0);
	if (!vd.validity.RowIsValid(idx)) throw InvalidInputException("pgq_b200: NULL csr_id / size argument");
	return ((int64_t *)vd.data)[idx];
}
static int32_t FirstInt(DataChunk &args, idx_t col) {
	VectorData vd;
	args.data[col].Orrify(args.size(), vd);
	auto idx = vd.sel->get_index(0);
	if (!vd.validity.RowIsValid(idx)) throw InvalidInputException("pgq_b200: NULL csr_id argument");
	return ((int32_t *)vd.data)[idx];
}

// dense copy of a BIGINT column + one validity byte per row
struct DenseCol {
	std::vector<int64_t> val;
	std::vector<uint8_t> valid;
	bool all_valid = true;
	void Load(Vector &v, idx_t count) {
		VectorData vd;
		v.Orrify(count, vd);
		val.resize(count);
		valid.resize(count);
		auto data = (int64_t *)vd.data;
		for (idx_t i = 0; i < count; i++) {
			auto idx = vd.sel->get_index(i);
			bool ok = vd.validity.RowIsValid(idx);
			valid[i] = ok;
			val[i] = ok ? data[idx] : 0;
			all_valid &= ok;
		}
	}
};

// ---- create_csr_vertex(INTEGER csr_id, BIGINT v_size, BIGINT vertex, BIGINT cnt) -> BIGINT (R2) ----
static void CreateCsrVertexFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	const int64_t v_size = FirstBigint(args, 1);
	DenseCol vid, cnt;
	vid.Load(args.data[2], count);
	cnt.Load(args.data[3], count);
	auto out = FlatVector::GetData<int64_t>(result);
	auto &validity = FlatVector::Validity(result);
	// rows with a NULL vertex id or count carry no degree: NULL out, not forwarded
	std::vector<int64_t> fv, fc;
	fv.reserve(count);
	fc.reserve(count);
	for (idx_t i = 0; i < count; i++) {
		if (!vid.valid[i] || !cnt.valid[i]) {
			validity.SetInvalid(i);
			out[i] = 0;
			continue;
		}
		fv.push_back(vid.val[i]);
		fc.push_back(cnt.val[i]);
		out[i] = cnt.val[i];
	}
	std::lock_guard<std::mutex> lock(st.mu);
	auto *pctx = st.Ctx();
	auto &slot = st.Slot(ContextOf(state), csr_id);
	if (!slot.vertex_started || slot.finalized) { // first row of a (new) CSR: CsrInitializeVertex
		PgqState::Check("create_csr_vertex", pgq_csr_create(pctx, slot.dev_id, v_size));
		slot.vertex_started = true;
		slot.edge_started = false;
		slot.finalized = false;
		slot.rows_added = 0;
	}
	int64_t sum = 0;
	PgqState::Check("create_csr_vertex",
	                pgq_csr_set_degrees(pctx, slot.dev_id, fv.data(), fc.data(), (int64_t)fv.size(), &sum));
}

// ---- create_csr_edge(INTEGER csr_id, BIGINT v_size, BIGINT num_edges, BIGINT src, BIGINT dst
//                      [, BIGINT edge_rowid]) -> INTEGER (R3) -------------------------------------
static void CreateCsrEdgeFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	const int64_t v_size = FirstBigint(args, 1);
	const int64_t num_edges = FirstBigint(args, 2);
	const bool with_eid = args.ColumnCount() >= 6;
	const bool with_w = args.ColumnCount() == 7; // later DuckPGQ: ..., edge_rowid, weight BIGINT|DOUBLE (R10)
	DenseCol src, dst, eid, wcol;
	src.Load(args.data[3], count);
	dst.Load(args.data[4], count);
	if (with_eid) eid.Load(args.data[5], count);
	if (with_w) wcol.Load(args.data[6], count); // BIGINT and DOUBLE are both 8-byte columns: raw copy
	if (!src.all_valid || !dst.all_valid || (with_eid && !eid.all_valid) || (with_w && !wcol.all_valid)) {
		throw InvalidInputException("pgq_b200: create_csr_edge got a NULL src / dst / edge id / weight");
	}
	auto out = FlatVector::GetData<int32_t>(result);
	for (idx_t i = 0; i < count; i++) out[i] = 1;
	std::lock_guard<std::mutex> lock(st.mu);
	auto *pctx = st.Ctx();
	auto &slot = st.Slot(ContextOf(state), csr_id);
	if (!slot.vertex_started || slot.finalized) { // edge rows without create_csr_vertex rows
		PgqState::Check("create_csr_edge", pgq_csr_create(pctx, slot.dev_id, v_size));
		slot.vertex_started = true;
		slot.edge_started = false;
		slot.finalized = false;
		slot.rows_added = 0;
	}
	if (slot.rows_added == 0) {
		slot.weights.clear();
		slot.wtype = !with_w ? 0 : args.data[6].GetType().id() == LogicalTypeId::DOUBLE ? PGQ_WEIGHT_F64 : PGQ_WEIGHT_I64;
	}
	if (with_w) {
		const uint64_t *raw = reinterpret_cast<const uint64_t *>(wcol.val.data());
		slot.weights.insert(slot.weights.end(), raw, raw + count);
	}
	if (!slot.edge_started) { // first row: CsrInitializeEdge
		PgqState::Check("create_csr_edge", pgq_csr_begin_edges(pctx, slot.dev_id, num_edges));
		slot.edge_started = true;
		slot.num_edges = num_edges;
	}
	// rows are appended in arrival order = the stable key of the CSR (deterministic with threads=1)
	PgqState::Check("create_csr_edge", pgq_csr_add_edges(pctx, slot.dev_id, src.val.data(), dst.val.data(),
	                                                     with_eid ? eid.val.data() : nullptr, (int64_t)count));
	slot.rows_added += (int64_t)count;
	if (slot.rows_added == slot.num_edges) st.Finish(slot); // last row
}

// ---- iterativelength / reachability (R4-R6) ----------------------------------------------------------
template <bool REACH>
static void LengthFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	// reachability's optional BOOLEAN variant flag sits at position 1 and does not change results
	const idx_t base = args.ColumnCount() == 5 ? 2 : 1;
	DenseCol src, dst;
	src.Load(args.data[base + 1], count);
	dst.Load(args.data[base + 2], count);
	std::vector<uint8_t> ov(count);
	auto &validity = FlatVector::Validity(result);
	std::lock_guard<std::mutex> lock(st.mu);
	auto *pctx = st.Ctx();
	auto &slot = st.ReadySlot(ContextOf(state), csr_id);
	if (REACH) {
		std::vector<uint8_t> ob(count);
		PgqState::Check("reachability", pgq_reachability(pctx, slot.dev_id, (int64_t)count, src.val.data(),
		                                                 dst.val.data(), src.valid.data(), dst.valid.data(),
		                                                 ob.data(), ov.data()));
		auto out = FlatVector::GetData<bool>(result);
		for (idx_t i = 0; i < count; i++) {
			out[i] = ob[i] != 0;
			if (!ov[i]) validity.SetInvalid(i);
		}
	} else {
		auto out = FlatVector::GetData<int64_t>(result);
		if (args.ColumnCount() == 6) {
			// iterativelength(csr_id, v_size, src, dst, lower, upper) of the later DuckPGQ code (R11): the
			// bounds are per-query constants, read from row 0 like csr_id / v_size
			const int64_t lower = FirstBigint(args, 4), upper = FirstBigint(args, 5);
			if (lower < 0 || upper < 0)
				throw InvalidInputException("pgq_b200: iterativelength bounds must be non-negative");
			PgqState::Check("iterativelength",
			                pgq_iterativelength_bounded(pctx, slot.dev_id, (int64_t)count, src.val.data(),
			                                            dst.val.data(), src.valid.data(), dst.valid.data(), lower,
			                                            upper, out, ov.data()));
		} else
		PgqState::Check("iterativelength", pgq_iterativelength(pctx, slot.dev_id, (int64_t)count, src.val.data(),
		                                                       dst.val.data(), src.valid.data(), dst.valid.data(),
		                                                       out, ov.data()));
		for (idx_t i = 0; i < count; i++) {
			if (!ov[i]) validity.SetInvalid(i);
		}
	}
}

// ---- cheapest_path_length -> BIGINT / DOUBLE (R10) --------------------------------------------------
// The result type of a scalar function is fixed at bind time, before the CSR (and its weight type)
// exists, so the two weight types get two names: cheapest_path_length (BIGINT weights) and
// cheapest_path_length_double (DOUBLE weights).
template <bool F64>
static void CheapestPathLengthFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	DenseCol src, dst;
	src.Load(args.data[2], count);
	dst.Load(args.data[3], count);
	std::vector<uint8_t> ov(count);
	auto &validity = FlatVector::Validity(result);
	std::lock_guard<std::mutex> lock(st.mu);
	auto *pctx = st.Ctx();
	auto &slot = st.ReadySlot(ContextOf(state), csr_id);
	int wtype = 0;
	PgqState::Check("cheapest_path_length", pgq_csr_weight_type(pctx, slot.dev_id, &wtype));
	if (wtype != (F64 ? PGQ_WEIGHT_F64 : PGQ_WEIGHT_I64)) {
		throw InvalidInputException(wtype == 0 ? "pgq_b200: this CSR has no weights (7-argument create_csr_edge)"
		                            : F64      ? "pgq_b200: BIGINT weights: use cheapest_path_length"
		                                       : "pgq_b200: DOUBLE weights: use cheapest_path_length_double");
	}
	void *out = F64 ? (void *)FlatVector::GetData<double>(result) : (void *)FlatVector::GetData<int64_t>(result);
	PgqState::Check("cheapest_path_length",
	                pgq_cheapest_path_length(pctx, slot.dev_id, (int64_t)count, src.val.data(), dst.val.data(),
	                                         src.valid.data(), dst.valid.data(), out, ov.data()));
	for (idx_t i = 0; i < count; i++) {
		if (!ov[i]) validity.SetInvalid(i);
	}
}

// ---- shortestpath -> LIST<BIGINT> (R7-R9) -------------------------------------------------------------
template <int MODE>
static void ShortestPathFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	DenseCol src, dst;
	src.Load(args.data[2], count);
	dst.Load(args.data[3], count);
	std::vector<uint8_t> ov(count);
	std::vector<pgq_list_entry_t> entries(count);
	int64_t *vals = nullptr;
	uint64_t total = 0;
	{
		std::lock_guard<std::mutex> lock(st.mu);
		auto *pctx = st.Ctx();
		auto &slot = st.ReadySlot(ContextOf(state), csr_id);
		PgqState::Check("shortestpath", pgq_shortestpath(pctx, slot.dev_id, (int64_t)count, src.val.data(),
		                                                 dst.val.data(), src.valid.data(), dst.valid.data(), MODE,
		                                                 entries.data(), ov.data(), &vals, &total));
	}
	auto list_data = FlatVector::GetData<list_entry_t>(result);
	auto &validity = FlatVector::Validity(result);
	for (idx_t i = 0; i < count; i++) {
		list_data[i].offset = entries[i].offset;
		list_data[i].length = ov[i] ? entries[i].length : 0;
		if (!ov[i]) validity.SetInvalid(i);
	}
	ListVector::Reserve(result, total);
	if (total) memcpy(FlatVector::GetData<int64_t>(ListVector::GetEntry(result)), vals, total * sizeof(int64_t));
	ListVector::SetListSize(result, total);
	pgq_free(vals);
}

// ---- delete_csr(INTEGER csr_id) -> BOOLEAN ------------------------------------------------------------
static void DeleteCsrFunction(DataChunk &args, ExpressionState &state, Vector &result) {
	const idx_t count = args.size();
	result.SetVectorType(VectorType::FLAT_VECTOR);
	if (count == 0) return;
	auto &st = State();
	const int32_t csr_id = FirstInt(args, 0);
	auto out = FlatVector::GetData<bool>(result);
	std::lock_guard<std::mutex> lock(st.mu);
	auto key = std::make_pair(ContextOf(state), csr_id);
	auto it = st.slots.find(key);
	bool existed = it != st.slots.end();
	if (existed) {
		if (it->second.vertex_started) pgq_csr_delete(st.Ctx(), it->second.dev_id);
		st.slots.erase(it);
	}
	for (idx_t i = 0; i < count; i++) out[i] = existed;
}

static void Register(ClientContext &context, Catalog &catalog, const string &name, vector<LogicalType> args,
                     LogicalType ret, scalar_function_t fn) {
	ScalarFunction f(name, move(args), move(ret), move(fn), /*has_side_effects=*/true, PgqBind);
	CreateScalarFunctionInfo info(f);
	info.on_conflict = OnCreateConflict::ERROR_ON_CONFLICT;
	catalog.CreateFunction(context, &info);
}

} // namespace duckdb

using namespace duckdb;

extern "C" {
DUCKDB_EXTENSION_API void pgq_b200_init(duckdb::DatabaseInstance &db) {
	Connection con(db);
	con.BeginTransaction();
	auto &context = *con.context;
	auto &catalog = Catalog::GetCatalog(context);
	const LogicalType I = LogicalType::INTEGER, B = LogicalType::BIGINT, BO = LogicalType::BOOLEAN;
	Register(context, catalog, "create_csr_vertex", {I, B, B, B}, B, CreateCsrVertexFunction);
	{
		// two overloads: with and without the edge rowid column
		ScalarFunctionSet set("create_csr_edge");
		set.AddFunction(ScalarFunction({I, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		// with an edge weight (later DuckPGQ, rule R10)
		set.AddFunction(ScalarFunction({I, B, B, B, B, B, B}, I, CreateCsrEdgeFunction, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B, LogicalType::DOUBLE}, I, CreateCsrEdgeFunction, true,
		                               PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	{
		// 4 arguments, and the later 6-argument form with lower / upper length bounds (f1, rule R11)
		ScalarFunctionSet set("iterativelength");
		set.AddFunction(ScalarFunction({I, B, B, B}, B, LengthFunction<false>, true, PgqBind));
		set.AddFunction(ScalarFunction({I, B, B, B, B, B}, B, LengthFunction<false>, true, PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	{
		ScalarFunctionSet set("reachability");
		set.AddFunction(ScalarFunction({I, B, B, B}, BO, LengthFunction<true>, true, PgqBind));
		set.AddFunction(ScalarFunction({I, BO, B, B, B}, BO, LengthFunction<true>, true, PgqBind));
		CreateScalarFunctionInfo info(set);
		catalog.CreateFunction(context, &info);
	}
	Register(context, catalog, "shortestpath", {I, B, B, B}, LogicalType::LIST(B), ShortestPathFunction<0>);
	// later DuckPGQ behaviour: [v, e, v, ..., v] with edge row ids interleaved
	Register(context, catalog, "shortestpath_with_edges", {I, B, B, B}, LogicalType::LIST(B),
	         ShortestPathFunction<1>);
	Register(context, catalog, "cheapest_path_length", {I, B, B, B}, B, CheapestPathLengthFunction<false>);
	Register(context, catalog, "cheapest_path_length_double", {I, B, B, B}, LogicalType::DOUBLE,
	         CheapestPathLengthFunction<true>);
	Register(context, catalog, "delete_csr", {I}, BO, DeleteCsrFunction);
	con.Commit();
}

DUCKDB_EXTENSION_API const char *pgq_b200_version() {
	return DuckDB::LibraryVersion();
}
}
