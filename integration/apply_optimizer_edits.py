#!/usr/bin/env python
"""apply_optimizer_edits.py SRC_DIR — the lance_optimizer.cpp part of the reference-side edits (see apply_edits.py):
match LOGICAL_TOP_N (DuckDB's TOP_N pass has fused LIMIT + ORDER BY before extension optimizers run), follow
the sort key through the projection that computes it, and carry scan-level table filters into the Lance predicate."""
import sys

p = sys.argv[1].rstrip("/") + "/lance_optimizer.cpp"
s = open(p).read()


def rep(old, new):
    global s
    assert s.count(old) >= 1, old[:80]
    s = s.replace(old, new, 1)


rep('''#include "duckdb/planner/operator/logical_projection.hpp"
''', '''#include "duckdb/planner/operator/logical_projection.hpp"
#include "duckdb/planner/operator/logical_top_n.hpp"
#include "duckdb/planner/filter/conjunction_filter.hpp"
#include "duckdb/planner/filter/constant_filter.hpp"
#include "duckdb/planner/filter/in_filter.hpp"
#include "duckdb/planner/filter/null_filter.hpp"
#include "duckdb/planner/filter/optional_filter.hpp"
''')

HELPER = '''// Convert a filter DuckDB's filter pushdown already moved into the scan (LogicalGet::table_filters) to a Lance
// predicate on column `col`.  `out` stays empty for filters that may be ignored (OPTIONAL_FILTER).
// Returns false for filters that cannot be expressed: the rewrite is then abandoned, never silently weakened.
static bool TableFilterToLancePredicate(const TableFilter &filter, const string &col, string &out) {
	switch (filter.filter_type) {
	case TableFilterType::CONSTANT_COMPARISON: {
		auto &cf = filter.Cast<ConstantFilter>();
		auto *op_str = ComparisonOperator(cf.comparison_type);
		string constant;
		if (!op_str || !FormatConstant(cf.constant, constant)) {
			return false;
		}
		out = col + " " + op_str + " " + constant;
		return true;
	}
	case TableFilterType::IS_NULL:
		out = col + " IS NULL";
		return true;
	case TableFilterType::IS_NOT_NULL:
		out = col + " IS NOT NULL";
		return true;
	case TableFilterType::IN_FILTER: {
		auto &in = filter.Cast<InFilter>();
		string list;
		for (auto &v : in.values) {
			string constant;
			if (!FormatConstant(v, constant)) {
				return false;
			}
			list += (list.empty() ? "" : ", ") + constant;
		}
		if (list.empty()) {
			return false;
		}
		out = col + " IN (" + list + ")";
		return true;
	}
	case TableFilterType::CONJUNCTION_AND:
	case TableFilterType::CONJUNCTION_OR: {
		bool is_and = filter.filter_type == TableFilterType::CONJUNCTION_AND;
		auto &children = is_and ? filter.Cast<ConjunctionAndFilter>().child_filters
		                        : filter.Cast<ConjunctionOrFilter>().child_filters;
		string joined;
		for (auto &child : children) {
			string part;
			if (!TableFilterToLancePredicate(*child, col, part)) {
				return false;
			}
			if (part.empty()) {
				if (!is_and) {
					return false; // an ignorable branch of an OR: keep it simple, no rewrite
				}
				continue;
			}
			joined += (joined.empty() ? "" : (is_and ? " AND " : " OR ")) + ("(" + part + ")");
		}
		out = joined.empty() ? "" : "(" + joined + ")";
		return true;
	}
	case TableFilterType::OPTIONAL_FILTER:
		out.clear(); // not required for correctness (zone-map style hints)
		return true;
	default:
		return false; // STRUCT_EXTRACT, DYNAMIC_FILTER, EXPRESSION_FILTER
	}
}

// ========================================
// Optimizer: detect ORDER BY array_distance(...) LIMIT k
// ========================================
'''
rep('''// ========================================
// Optimizer: detect ORDER BY array_distance(...) LIMIT k
// ========================================
''', HELPER)

a = s.index('		// Match: LogicalLimit -> LogicalOrder -> ... -> LogicalGet')
b = s.index('		auto &func_expr = order_expr->Cast<BoundFunctionExpression>();')
NEW_HEAD = '''		// Two plan shapes carry "ORDER BY <distance> LIMIT k":
		//  (A) LOGICAL_TOP_N — what an extension optimizer actually sees: DuckDB's TOP_N pass fuses LIMIT + ORDER BY
		//      before the extension hooks run (optimizer.cpp: built-in passes first; topn_optimizer.cpp).
		//  (B) LOGICAL_LIMIT -> LOGICAL_ORDER_BY — kept for the cases TopN declines (large k relative to the table).
		LogicalOperator *sort_node = nullptr; // the node whose child is the subtree to replace
		vector<BoundOrderByNode> *orders = nullptr;
		idx_t limit_val = 0;
		if (op->type == LogicalOperatorType::LOGICAL_TOP_N) {
			auto &topn = op->Cast<LogicalTopN>();
			if (topn.offset > 0) {
				return; // index scan returns top-k, can't skip rows
			}
			limit_val = topn.limit;
			sort_node = op.get();
			orders = &topn.orders;
		} else if (op->type == LogicalOperatorType::LOGICAL_LIMIT) {
			auto &limit_op = op->Cast<LogicalLimit>();

			// Must be a constant limit (not percentage or expression)
			if (limit_op.limit_val.Type() != LimitNodeType::CONSTANT_VALUE) {
				return;
			}
			limit_val = limit_op.limit_val.GetConstantValue();

			// Bail if OFFSET is present — index scan returns top-k, can't skip rows
			if (limit_op.offset_val.Type() != LimitNodeType::UNSET &&
			    !(limit_op.offset_val.Type() == LimitNodeType::CONSTANT_VALUE &&
			      limit_op.offset_val.GetConstantValue() == 0)) {
				return;
			}

			if (op->children.empty() || op->children[0]->type != LogicalOperatorType::LOGICAL_ORDER_BY) {
				return;
			}
			auto &order_op = op->children[0]->Cast<LogicalOrder>();
			sort_node = &order_op;
			orders = &order_op.orders;
		} else {
			return;
		}
		if (limit_val == 0 || sort_node->children.empty()) {
			return;
		}

		if (orders->size() != 1) {
			return;
		}

		// Only rewrite ASC ordering — DESC wants farthest vectors, not nearest
		if ((*orders)[0].type == OrderType::DESCENDING) {
			return;
		}

		// The sort key is either the distance call itself or — the usual case, because the binder evaluates ORDER BY
		// expressions in the projection below the sort — a column reference into that projection.
		Expression *order_expr = (*orders)[0].expression.get();
		LogicalProjection *key_proj = nullptr;
		if (order_expr->type == ExpressionType::BOUND_COLUMN_REF) {
			auto &key_ref = order_expr->Cast<BoundColumnRefExpression>();
			auto *below = sort_node->children[0].get();
			if (below->type != LogicalOperatorType::LOGICAL_PROJECTION) {
				return;
			}
			auto &proj = below->Cast<LogicalProjection>();
			if (key_ref.binding.table_index != proj.table_index ||
			    key_ref.binding.column_index >= proj.expressions.size() || proj.children.empty()) {
				return;
			}
			key_proj = &proj;
			order_expr = proj.expressions[key_ref.binding.column_index].get();
		}

		if (!IsArrayDistanceFunction(*order_expr)) {
			return;
		}

'''
s = s[:a] + NEW_HEAD + s[b:]

rep('''		auto *walk = order_op.children[0].get();''',
    '''		// With a key projection (see above) that projection sits between the sort and this chain and is kept.
		auto &subtree = key_proj ? key_proj->children[0] : sort_node->children[0];
		auto *walk = subtree.get();''')

rep('''		// Try to push filter predicates to Lance
		bool filter_fully_pushed = false;''',
    '''		// Filters that DuckDB's filter pushdown already moved into the scan must travel too: the replacement scan
		// does not evaluate table_filters.  All of them convert, or there is no rewrite.
		for (auto &entry : get.table_filters.filters) {
			auto &table_columns = duck_table.GetColumns();
			if (entry.first >= table_columns.PhysicalColumnCount()) {
				return;
			}
			const string col_name = table_columns.GetColumn(PhysicalIndex(entry.first)).GetName();
			string pred_sql;
			if (!TableFilterToLancePredicate(*entry.second, col_name, pred_sql)) {
				return;
			}
			if (pred_sql.empty()) {
				continue;
			}
			if (!bind_data->predicate.empty()) {
				bind_data->predicate += " AND ";
			}
			bind_data->predicate += pred_sql;
		}

		// Try to push filter predicates to Lance
		bool filter_fully_pushed = false;''')

# DuckDB 1.4's late-materialization pass (optimizer.cpp: runs right after TOP_N) turns "TOP_N over a scan" into a
# semi-join on the row id whose right side is TOP_N -> PROJ -> GET(.., rowid): the replacement GET must know the
# scan's virtual columns, or binding the row id column throws "Failed to find referenced virtual column".
rep('''		    make_uniq<LogicalGet>(get.table_index, scan_func, std::move(bind_data), get.returned_types, get.names);''',
    '''		    make_uniq<LogicalGet>(get.table_index, scan_func, std::move(bind_data), get.returned_types, get.names,
		                          get.virtual_columns);''')

# The scan already emits exactly the requested columns (input.column_ids): declare it, so that virtual columns (rowid)
# are legal on it (plan_get.cpp: "Virtual columns require projection pushdown"), and honour projection_ids when the
# original scan had filter-only columns pruned.
rep('''		TableFunction scan_func("lance_index_scan", {}, LanceIndexScanScan, LanceIndexScanBind, LanceIndexScanInit);
''', '''		TableFunction scan_func("lance_index_scan", {}, LanceIndexScanScan, LanceIndexScanBind, LanceIndexScanInit);
		scan_func.projection_pushdown = true;
''')
rep('''	for (auto &col_id : input.column_ids) {
		state->storage_ids.push_back(StorageIndex(col_id));
	}
''', '''	if (input.projection_ids.empty()) {
		for (auto &col_id : input.column_ids) {
			state->storage_ids.push_back(StorageIndex(col_id));
		}
	} else {
		for (auto &proj_id : input.projection_ids) {
			state->storage_ids.push_back(StorageIndex(input.column_ids[proj_id]));
		}
	}
''')

a = s.index('		// Replace the LIMIT → ORDER → [FILTER →] [PROJ →] GET subtree')
b = s.index('void RegisterLanceOptimizer')
NEW_TAIL = '''		// Replace the [FILTER ->] [PROJ ->] GET chain below the sort (below the key projection when there is one)
		unique_ptr<LogicalOperator> replacement;
		if (filter_ptr && !filter_fully_pushed) {
			// Partial pushdown: keep FILTER, replace its child subtree with new GET
			if (has_projection) {
				// FILTER -> PROJ -> GET  =>  FILTER -> PROJ -> new GET
				auto *proj_node = filter_ptr->children[0].get();
				proj_node->children[0] = std::move(new_get);
			} else {
				// FILTER -> GET  =>  FILTER -> new GET
				filter_ptr->children[0] = std::move(new_get);
			}
			replacement = std::move(subtree);
		} else if (has_projection) {
			// [FILTER ->] PROJ -> GET  =>  PROJ -> new GET
			unique_ptr<LogicalOperator> proj;
			if (filter_ptr) {
				// filter fully pushed: FILTER -> PROJ -> GET
				proj = std::move(filter_ptr->children[0]);
			} else {
				proj = std::move(subtree);
			}
			proj->children[0] = std::move(new_get);
			replacement = std::move(proj);
		} else {
			// [FILTER ->] GET  =>  new GET
			replacement = std::move(new_get);
		}

		// The index scan emits rows in ascending distance, so the sort (and the limit) go away.  A key projection
		// stays: it still produces the columns the operators above reference.
		if (key_proj) {
			key_proj->children[0] = std::move(replacement);
			op = std::move(sort_node->children[0]);
		} else {
			op = std::move(replacement);
		}
	}
};

'''
s = s[:a] + NEW_TAIL + s[b:]
open(p, "w").write(s)
