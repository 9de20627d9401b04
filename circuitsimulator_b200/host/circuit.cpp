// circuit.cpp — Circuit graph and MNA numbering (reference behaviour: src/circuit.cpp).
#include "csim/circuit.hpp"

#include <cmath>
#include <iostream>

int Circuit::getOrCreateNode(const std::string& name) {
    auto found = nodeNameToId.find(name);
    if (found != nodeNameToId.end()) return found->second;
    const int id = static_cast<int>(nodes.size());
    nodes.emplace_back(id, name);
    nodeNameToId.emplace(name, id);
    return id;
}

int Circuit::numNodeEquations() const {
    int n = 0;
    for (const Node& nd : nodes) n += isGroundName(nd.name) ? 0 : 1;
    return n;
}

// one extra unknown (branch current) per voltage source and per inductor
int Circuit::numVoltageBranches() const {
    int n = 0;
    for (const auto& e : elements) {
        const int ty = e->abiType();
        if (ty == 3 /*V*/ || ty == 2 /*L*/) ++n;
    }
    return n;
}

int Circuit::numUnknowns() const { return numNodeEquations() + numVoltageBranches(); }

// reference: src/circuit.cpp:42-61 — node rows in creation (first appearance) order with
// ground = -1, then branch rows for V and L in element order.
void Circuit::assignEquationIndices() {
    int next = 0;
    for (Node& nd : nodes) nd.eqIndex = isGroundName(nd.name) ? -1 : next++;
    for (auto& e : elements) {
        if (auto* v = dynamic_cast<VoltageSource*>(e.get())) v->setBranchEqIndex(next++);
        else if (auto* l = dynamic_cast<Inductor*>(e.get())) l->setBranchEqIndex(next++);
    }
}

void Circuit::attach(const std::shared_ptr<Element>& e, std::initializer_list<int> ids) {
    const int idx = static_cast<int>(elements.size());
    elements.push_back(e);
    for (int id : ids) nodes[id].attachedElements.push_back(idx);
}

void Circuit::addResistor(const std::string& name, const std::string& n1, const std::string& n2, double value) {
    const int a = getOrCreateNode(n1), b = getOrCreateNode(n2);
    attach(std::make_shared<Resistor>(name, a, b, value), {a, b});
}
void Circuit::addCapacitor(const std::string& name, const std::string& n1, const std::string& n2, double value) {
    const int a = getOrCreateNode(n1), b = getOrCreateNode(n2);
    attach(std::make_shared<CapacitorElement>(name, a, b, value), {a, b});
}
void Circuit::addInductor(const std::string& name, const std::string& n1, const std::string& n2, double value) {
    const int a = getOrCreateNode(n1), b = getOrCreateNode(n2);
    attach(std::make_shared<Inductor>(name, a, b, value), {a, b});
}
void Circuit::addCurrentSource(const std::string& name, const std::string& np, const std::string& nm,
                               const SourceSpec& spec) {
    const int p = getOrCreateNode(np), m = getOrCreateNode(nm);
    attach(std::make_shared<CurrentSource>(name, p, m, spec), {p, m});
}
void Circuit::addVoltageSource(const std::string& name, const std::string& np, const std::string& nm,
                               const SourceSpec& spec) {
    const int p = getOrCreateNode(np), m = getOrCreateNode(nm);
    attach(std::make_shared<VoltageSource>(name, p, m, spec), {p, m});
}

// reference: src/circuit.cpp:128-168 — K = MU*COX*(W/L), Vth = |VT|, bulk tied to node "0",
// channel type from the model (sign of VT, decided when the .MODEL card is read).
void Circuit::addMosfet(const std::string& name, const std::string& nd, const std::string& ng, const std::string& ns,
                        const std::string& modelId, double W, double L) {
    const MosModel* m = findMosModel(modelId);
    if (!m) {
        std::cerr << "Unknown MOS model: " << modelId << "\n";
        return;
    }
    const int d = getOrCreateNode(nd), g = getOrCreateNode(ng), s = getOrCreateNode(ns), b = getOrCreateNode("0");
    const double K = m->MU * m->COX * (W / L);
    const double vth = std::abs(m->VT);
    std::shared_ptr<Element> e;
    if (m->isP) e = std::make_shared<PMosElement>(name, d, g, s, b, vth, K, m->LAMBDA, m->CJO);
    else e = std::make_shared<NMosElement>(name, d, g, s, b, vth, K, m->LAMBDA, m->CJO);
    attach(e, {d, g, s, b});
}

void Circuit::addMosModel(const MosModel& m) { mosModels[m.name] = m; }

const MosModel* Circuit::findMosModel(const std::string& id) const {
    auto it = mosModels.find(id);
    return it == mosModels.end() ? nullptr : &it->second;
}

void Circuit::printConnectivity() const {
    std::cout << "========== nodes and attached elements ==========\n";
    for (const Node& nd : nodes) {
        std::cout << "Node " << nd.name << " (id=" << nd.id << ", eqIndex=" << nd.eqIndex << "): ";
        for (int ei : nd.attachedElements) std::cout << elements[ei]->getName() << " ";
        std::cout << "\n";
    }
}
