// Exact replay of libstdc++'s std::sort on an index array with the comparator key[a] < key[b].
//
// Why: fillstats (src/aQueryFasta_thread.cpp:319-327) orders a pair's unique k-mers with an UNSTABLE
// std::sort by "number of mapped loci"; almost every k-mer ties, and find_matching_locus
// (cpp:364-422) stops early, so the tie permutation decides top.fc/top.rc and, for reads split
// between loci, the locus itself (SURVEY.md F4 trap 5).  The sequence of comparisons and moves below
// is the one GCC 13's <bits/stl_algo.h> performs (__GLIBCXX__ 20240904): __sort :1942,
// __introsort_loop :1918 (threshold 16, depth 2*floor(log2 n)), __move_median_to_first :85,
// __unguarded_partition :1871, __partial_sort/heap fallback :1905 + <bits/stl_heap.h>:135-425,
// __final_insertion_sort :1854.  It is written iteratively (explicit stack) for one GPU lane over
// shared-memory arrays; tests/test_introsort.py replays it against the host's std::sort.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define DBT_SORT_HD __host__ __device__
#else
#define DBT_SORT_HD
#endif
#ifndef DBT_SORT_ON_HEAPSORT
#define DBT_SORT_ON_HEAPSORT()      // test hook: counts entries into the heap-sort fallback
#endif

namespace dbt {

template <typename IdxT, typename KeyT>
struct IntroSort {
	IdxT* a;            // index array being permuted
	const KeyT* key;    // key[a[i]] is the sort key

	DBT_SORT_HD bool lt(IdxT x, IdxT y) const { return key[x] < key[y]; }
	DBT_SORT_HD void swp(int i, int j) { IdxT t = a[i]; a[i] = a[j]; a[j] = t; }

	DBT_SORT_HD void unguarded_linear_insert(int last) {
		IdxT val = a[last];
		int next = last - 1;
		while (lt(val, a[next])) { a[last] = a[next]; last = next; --next; }
		a[last] = val;
	}
	DBT_SORT_HD void insertion_sort(int first, int last) {
		if (first == last) return;
		for (int i = first + 1; i != last; ++i) {
			if (lt(a[i], a[first])) {
				IdxT val = a[i];
				for (int j = i; j > first; --j) a[j] = a[j - 1];   // move_backward
				a[first] = val;
			} else unguarded_linear_insert(i);
		}
	}
	DBT_SORT_HD void push_heap(int first, int hole, int top, IdxT value) {
		int parent = (hole - 1) / 2;
		while (hole > top && lt(a[first + parent], value)) { a[first + hole] = a[first + parent]; hole = parent; parent = (hole - 1) / 2; }
		a[first + hole] = value;
	}
	DBT_SORT_HD void adjust_heap(int first, int hole, int len, IdxT value) {
		const int top = hole;
		int child = hole;
		while (child < (len - 1) / 2) {
			child = 2 * (child + 1);
			if (lt(a[first + child], a[first + child - 1])) child--;
			a[first + hole] = a[first + child];
			hole = child;
		}
		if ((len & 1) == 0 && child == (len - 2) / 2) {
			child = 2 * (child + 1);
			a[first + hole] = a[first + child - 1];
			hole = child - 1;
		}
		push_heap(first, hole, top, value);
	}
	DBT_SORT_HD void heap_sort(int first, int last) {          // __partial_sort(first, last, last)
		int len = last - first;
		if (len >= 2) {
			int parent = (len - 2) / 2;
			for (;;) { adjust_heap(first, parent, len, a[first + parent]); if (parent == 0) break; parent--; }
		}
		while (last - first > 1) {
			--last;
			IdxT value = a[last];
			a[last] = a[first];
			adjust_heap(first, 0, last - first, value);
		}
	}
	DBT_SORT_HD void move_median_to_first(int result, int x, int y, int z) {
		if (lt(a[x], a[y])) {
			if (lt(a[y], a[z])) swp(result, y);
			else if (lt(a[x], a[z])) swp(result, z);
			else swp(result, x);
		} else if (lt(a[x], a[z])) swp(result, x);
		else if (lt(a[y], a[z])) swp(result, z);
		else swp(result, y);
	}
	DBT_SORT_HD int unguarded_partition(int first, int last, int pivot) {
		for (;;) {
			while (lt(a[first], a[pivot])) ++first;
			--last;
			while (lt(a[pivot], a[last])) --last;
			if (!(first < last)) return first;
			swp(first, last);
			++first;
		}
	}
	DBT_SORT_HD void sort(int n) {
		if (n <= 0) return;
		int lg = 0;
		for (int m = n; m > 1; m >>= 1) ++lg;
		// __introsort_loop(first, last, depth): the recursive call handles [cut, last) FIRST, then the
		// loop continues on [first, cut).  Explicit stack of pending [first, cut) ranges.
		struct Frame { int first, last, depth; };
		Frame stack[48];
		int sp = 0;
		int first = 0, last = n, depth = 2 * lg;
		for (;;) {
			if (last - first > 16) {
				if (depth == 0) { DBT_SORT_ON_HEAPSORT(); heap_sort(first, last); }
				else {
					--depth;
					int mid = first + (last - first) / 2;
					move_median_to_first(first, first + 1, mid, last - 1);
					int cut = unguarded_partition(first + 1, last, first);
					// recurse into [cut,last) now; remember [first,cut) for afterwards
					stack[sp].first = first; stack[sp].last = cut; stack[sp].depth = depth; ++sp;
					first = cut;
					continue;
				}
			}
			if (sp == 0) break;
			--sp;
			first = stack[sp].first; last = stack[sp].last; depth = stack[sp].depth;
		}
		if (n > 16) {
			insertion_sort(0, 16);
			for (int i = 16; i != n; ++i) unguarded_linear_insert(i);
		} else insertion_sort(0, n);
	}
};

}  // namespace dbt
