// plan.cpp — symbolic replay of the reference's assembly order (see plan.hpp).
#include "plan.hpp"

#include <algorithm>
#include <map>
#include <set>
#include <stdexcept>
#include <tuple>

namespace csim {
namespace {

struct Builder {
    ModeProgram& mp;
    std::map<std::pair<int, int>, int> static_memo;
    explicit Builder(ModeProgram& m) : mp(m) {}

    int new_value() { return mp.n_values++; }
    int stat(int kind, int param) {
        auto key = std::make_pair(kind, param);
        auto it = static_memo.find(key);
        if (it != static_memo.end()) return it->second;
        int v = new_value();
        mp.static_ops.push_back({kind, param, v, 0});
        static_memo[key] = v;
        return v;
    }
    int step(int kind, int g, int a, int b) {
        int v = new_value();
        mp.step_ops.push_back({kind, g, a, b, v, 0, 0, 0});
        return v;
    }
    void g(int row, int col, int v, bool sub) {
        if (row < 0 || col < 0) return;  // ground rows/columns are skipped (eq >= 0 guards in every stamp)
        mp.contribs.push_back({row, col, v, sub ? 1 : 0});
    }
    void rhs(int row, int v, bool sub) {
        if (row < 0) return;
        mp.contribs.push_back({row, -1, v, sub ? 1 : 0});
    }
    // Resistor::stamp / stampCapBE conductance block: (a,a)+ (b,b)+ (a,b)- (b,a)-
    void conductance(int a, int b, int v) {
        g(a, a, v, false);
        g(b, b, v, false);
        if (a >= 0 && b >= 0) {
            g(a, b, v, true);
            g(b, a, v, true);
        }
    }
    // VoltageSource / Inductor incidence block: (p,k)+ (m,k)- (k,p)+ (k,m)-
    void branch(int p, int m, int k, int one) {
        g(p, k, one, false);
        g(m, k, one, true);
        g(k, p, one, false);
        g(k, m, one, true);
    }
    // stampCapBE, src/tanalisis.cpp:59-80
    void cap_be(int a, int b, int gv) {
        int hist = step(TK_CAP_HIST, gv, a, b);
        conductance(a, b, gv);
        rhs(a, hist, true);   // I(eq1) -= I_hist
        rhs(b, hist, false);  // I(eq2) += I_hist
    }
    // MosfetBase::stamp rows D and S, src/element.cpp:289-304
    void mos_channel(int base, int D, int G, int S) {
        const int gd = base, gg = base + 1, gs = base + 2, cst = base + 3;
        if (D >= 0) {
            g(D, D, gd, false);
            g(D, G, gg, false);
            g(D, S, gs, false);
            rhs(D, cst, true);
        }
        if (S >= 0) {
            g(S, D, gd, true);  // += -gd
            g(S, G, gg, true);
            g(S, S, gs, true);
            rhs(S, cst, false);
        }
    }
};

void finalize(ModeProgram& mp, int N) {
    std::set<std::pair<int, int>> pat;
    for (auto& c : mp.contribs)
        if (c.col >= 0) pat.insert({c.row, c.col});
    mp.rowptr.assign(N + 1, 0);
    mp.colidx.clear();
    std::map<std::pair<int, int>, int> slot_of;
    for (auto& rc : pat) {
        slot_of[rc] = (int)mp.colidx.size();
        mp.colidx.push_back(rc.second);
        mp.rowptr[rc.first + 1]++;
    }
    for (int i = 0; i < N; ++i) mp.rowptr[i + 1] += mp.rowptr[i];
    const int nnz = (int)mp.colidx.size();
    std::vector<std::vector<uint32_t>> slots(nnz + N);
    std::vector<std::vector<uint32_t>> rows(N);
    for (auto& c : mp.contribs) {
        int slot = c.col >= 0 ? slot_of[{c.row, c.col}] : nnz + c.row;
        slots[slot].push_back(contrib_word(0, c.sub != 0, c.value));
        rows[c.row].push_back(contrib_word(c.col >= 0 ? c.col : kRhsCol, c.sub != 0, c.value));
    }
    mp.slot_ptr.assign(1, 0);
    mp.slot_entries.clear();
    for (auto& s : slots) {
        mp.slot_entries.insert(mp.slot_entries.end(), s.begin(), s.end());
        mp.slot_ptr.push_back((int)mp.slot_entries.size());
    }
    mp.row_len_max = 0;
    for (auto& r : rows) mp.row_len_max = std::max<int>(mp.row_len_max, (int)r.size());
    mp.row_prog.assign((size_t)mp.row_len_max * 32, 0xFFFFFFFFu);
    for (int i = 0; i < N && i < 32; ++i)
        for (size_t c = 0; c < rows[i].size(); ++c) mp.row_prog[c * 32 + i] = rows[i][c];
}

}  // namespace

HostPlan build_host_plan(const csim_topology& t, const csim_constants* consts) {
    HostPlan hp;
    if (t.n_unknowns <= 0 || t.n_elems < 0 || t.n_node_eq < 0 || t.n_node_eq > t.n_unknowns)
        throw std::runtime_error("topology: bad dimensions");
    if (!t.elem_type || !t.elem_flags || !t.elem_nodes || !t.elem_branch || !t.elem_param)
        throw std::runtime_error("topology: null array");
    hp.N = t.n_unknowns;
    hp.n_node_eq = t.n_node_eq;
    hp.n_params = t.n_params;
    if (consts) hp.consts = *consts;
    else csim_default_constants(&hp.consts);
    const int N = hp.N;
    hp.elem_type.assign(t.elem_type, t.elem_type + t.n_elems);
    hp.elem_flags.assign(t.elem_flags, t.elem_flags + t.n_elems);
    hp.elem_nodes.assign(t.elem_nodes, t.elem_nodes + 4 * (size_t)t.n_elems);
    hp.elem_branch.assign(t.elem_branch, t.elem_branch + t.n_elems);
    hp.elem_param.assign(t.elem_param, t.elem_param + t.n_elems);

    static const int kParams[6] = {CSIM_PARAMS_R, CSIM_PARAMS_C, CSIM_PARAMS_L, CSIM_PARAMS_SRC, CSIM_PARAMS_SRC,
                                   CSIM_PARAMS_MOS};
    for (int e = 0; e < t.n_elems; ++e) {
        int ty = t.elem_type[e];
        if (ty < 0 || ty > CSIM_ELEM_MOS) throw std::runtime_error("topology: unknown element type");
        if (t.elem_param[e] < 0 || t.elem_param[e] + kParams[ty] > t.n_params)
            throw std::runtime_error("topology: parameter offset out of range");
        for (int k = 0; k < 4; ++k) {
            int eq = t.elem_nodes[4 * e + k];
            if (eq < -1 || eq >= N) throw std::runtime_error("topology: node equation out of range");
        }
        bool needs_branch = (ty == CSIM_ELEM_V || ty == CSIM_ELEM_L);
        int k = t.elem_branch[e];
        if (needs_branch && (k < t.n_node_eq || k >= N))
            throw std::runtime_error("topology: V/L element without a valid branch equation "
                                     "(assignEquationIndices not called?)");
        if (ty == CSIM_ELEM_V || ty == CSIM_ELEM_I) hp.n_sources++;
        if (ty == CSIM_ELEM_MOS) {
            hp.n_mos++;
            hp.has_mos = true;
        }
    }

    for (int mode = 0; mode < 2; ++mode) {
        ModeProgram& mp = hp.mode[mode];
        Builder b(mp);
        const int vONE = b.stat(SK_ONE, -1);
        const int vGMIN = b.stat(SK_GMIN, -1);
        // independent-source values, element order (both modes)
        std::vector<int> src_val_of(t.n_elems, -1);
        {
            int s = 0;
            for (int e = 0; e < t.n_elems; ++e) {
                int ty = t.elem_type[e];
                if (ty != CSIM_ELEM_V && ty != CSIM_ELEM_I) continue;
                int v = b.new_value();
                mp.src_ops.push_back({v, t.elem_param[e], (t.elem_flags[e] & CSIM_FLAG_SRC_SIN) ? 1 : 0, s++});
                src_val_of[e] = v;
            }
        }
        // per-MOSFET linearisation values, element order (both modes)
        std::vector<int> mos_base_of(t.n_elems, -1);
        for (int e = 0; e < t.n_elems; ++e)
            if (t.elem_type[e] == CSIM_ELEM_MOS) {
                int base = b.new_value();
                b.new_value();
                b.new_value();
                b.new_value();
                const int32_t* nd = t.elem_nodes + 4 * e;
                mp.mos_ops.push_back(
                    {base, t.elem_param[e], nd[0], nd[1], nd[2], (t.elem_flags[e] & CSIM_FLAG_PMOS) ? 1 : 0, 0, 0});
                mos_base_of[e] = base;
            }

        auto stamp_rvi = [&](int e) {  // Element::stamp of R / I / V, src/element.cpp:9-32,34-66,83-123
            int ty = t.elem_type[e];
            const int32_t* nd = t.elem_nodes + 4 * e;
            if (ty == CSIM_ELEM_R) {
                b.conductance(nd[0], nd[1], b.stat(SK_RES, t.elem_param[e]));
            } else if (ty == CSIM_ELEM_I) {
                b.rhs(nd[0], src_val_of[e], true);
                b.rhs(nd[1], src_val_of[e], false);
            } else if (ty == CSIM_ELEM_V) {
                b.branch(nd[0], nd[1], t.elem_branch[e], vONE);
                b.rhs(t.elem_branch[e], src_val_of[e], false);
            }
        };

        if (mode == CSIM_MODE_DC) {
            // src/dcanalysis.cpp:126-130: all elements in netlist order, then gmin
            for (int e = 0; e < t.n_elems; ++e) {
                int ty = t.elem_type[e];
                const int32_t* nd = t.elem_nodes + 4 * e;
                if (ty == CSIM_ELEM_R || ty == CSIM_ELEM_I || ty == CSIM_ELEM_V) stamp_rvi(e);
                else if (ty == CSIM_ELEM_L) b.branch(nd[0], nd[1], t.elem_branch[e], vONE);  // src/element.cpp:156-178
                else if (ty == CSIM_ELEM_MOS) b.mos_channel(mos_base_of[e], nd[0], nd[1], nd[2]);
                // capacitors: open circuit (include/element.hpp:103-108)
            }
        } else {
            // src/tanalisis.cpp:268-353
            for (int e = 0; e < t.n_elems; ++e) {  // 1) R, V, I
                int ty = t.elem_type[e];
                if (ty == CSIM_ELEM_R || ty == CSIM_ELEM_I || ty == CSIM_ELEM_V) stamp_rvi(e);
            }
            for (int e = 0; e < t.n_elems; ++e)  // 2) MOS channel
                if (t.elem_type[e] == CSIM_ELEM_MOS) {
                    const int32_t* nd = t.elem_nodes + 4 * e;
                    b.mos_channel(mos_base_of[e], nd[0], nd[1], nd[2]);
                }
            for (int e = 0; e < t.n_elems; ++e)  // 3) explicit capacitors
                if (t.elem_type[e] == CSIM_ELEM_C) {
                    const int32_t* nd = t.elem_nodes + 4 * e;
                    b.cap_be(nd[0], nd[1], b.stat(SK_CAP, t.elem_param[e]));
                }
            for (int e = 0; e < t.n_elems; ++e)  // 4) inductors
                if (t.elem_type[e] == CSIM_ELEM_L) {
                    const int32_t* nd = t.elem_nodes + 4 * e;
                    int k = t.elem_branch[e];
                    int one = b.stat(SK_IND_ONE, t.elem_param[e]);
                    int nreq = b.stat(SK_IND_NREQ, t.elem_param[e]);
                    int hist = b.step(TK_IND_HIST, nreq, k, -1);
                    b.branch(nd[0], nd[1], k, one);
                    b.g(k, k, nreq, false);  // G(k,k) += -R_eq
                    b.rhs(k, hist, false);   // I(k)   += V_hist
                }
            for (int e = 0; e < t.n_elems; ++e)  // 5) MOS parasitic capacitors
                if (t.elem_type[e] == CSIM_ELEM_MOS) {
                    const int32_t* nd = t.elem_nodes + 4 * e;
                    int D = nd[0], G = nd[1], S = nd[2], B = nd[3];
                    int gh = b.stat(SK_CAPHALF, t.elem_param[e] + 3);
                    int gj = b.stat(SK_CAP, t.elem_param[e] + 3);
                    b.cap_be(G, S, gh);
                    b.cap_be(G, D, gh);
                    b.cap_be(S, B, gj);
                    b.cap_be(D, B, gj);
                }
        }
        for (int eq = 0; eq < t.n_node_eq; ++eq) b.g(eq, eq, vGMIN, false);  // stampGlobalGmin
        finalize(mp, N);
    }
    return hp;
}

}  // namespace csim
