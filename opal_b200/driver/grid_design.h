// grid_design.h -- partition of the global grid into x-slabs: GridDesign::{unbalanced, balanced}, grid/mod.rs:122-206.
#pragma once
#include <algorithm>
#include <functional>
#include <stdexcept>
#include <vector>

namespace opal_driver {

struct GridDesign {
    int numtasks = 1;
    double xmin = 0.0, dx = 0.0;
    std::vector<size_t> nx;          // cells per slab
    std::vector<double> offset;      // distance of each slab's left edge from xmin

    static std::vector<double> offsets(const std::vector<size_t>& nx, double dx) {     // the `scan` of mod.rs:131-138, 183-190
        std::vector<double> off;
        size_t rt = 0;
        for (size_t n : nx) { off.push_back((double)rt * dx); rt += n; }
        return off;
    }

    // even split; nx mod numtasks cells are dropped (quirk Q5, mod.rs:128)
    static GridDesign unbalanced(int numtasks, size_t size, double xmin, double dx, size_t min_subsize) {
        GridDesign d;
        d.numtasks = numtasks; d.xmin = xmin; d.dx = dx;
        const size_t subsize = std::max(size / (size_t)numtasks, min_subsize);
        d.nx.assign((size_t)numtasks, subsize);
        d.offset = offsets(d.nx, dx);
        return d;
    }

    // equal integral of the electron density per slab (mod.rs:157-206)
    static GridDesign balanced(int numtasks, size_t size, double xmin, double dx, size_t min_subsize, const std::function<double(double)>& ne) {
        if (numtasks == 0) return unbalanced(numtasks, size, xmin, dx, min_subsize);
        if (size < min_subsize) throw std::runtime_error("grid smaller than the minimum slab size");
        std::vector<double> cumsum;
        double total = 0.0;
        for (size_t n = 0; n < size - min_subsize; n++) {                  // the last slab keeps min_subsize cells
            const double x = xmin + (double)n * dx;
            total += dx * ne(x);
            cumsum.push_back(total);
        }
        if (cumsum.empty()) throw std::runtime_error("grid too small to balance");
        const double target = cumsum.back() / (double)numtasks;
        std::vector<size_t> ncells;
        size_t start = 0;
        for (int p = 1; p < numtasks; p++) {
            size_t i = 0;
            bool found = false;
            for (size_t k = start + min_subsize; k < cumsum.size(); k++, i++)
                if (cumsum[k] >= target * (double)p) { found = true; break; }
            if (!found) throw std::runtime_error("load balancing failed (the reference panics here, mod.rs:177)");
            ncells.push_back(i + min_subsize);
            start += i + min_subsize;
        }
        size_t ndone = 0;
        for (size_t n : ncells) ndone += n;
        ncells.push_back(size - ndone);
        GridDesign d;
        d.numtasks = numtasks; d.xmin = xmin; d.dx = dx;
        d.nx = ncells;
        d.offset = offsets(d.nx, dx);
        return d;
    }
};

}  // namespace opal_driver
