// circuit.hpp — netlist graph: nodes, elements, MOS model table, MNA numbering.
// Public surface of the reference's include/circuit.hpp.  assignEquationIndices() defines
// the equation order (non-ground nodes by first appearance, then one branch row per V and L
// in element order) and, with the element order, the MNA pattern and accumulation order.
#pragma once
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "element.hpp"
#include "utils.hpp"

struct Node {
    int id;
    std::string name;
    int eqIndex;                        // MNA row/column, -1 for ground
    std::vector<int> attachedElements;  // indices into Circuit::elements
    Node(int i, const std::string& n) : id(i), name(n), eqIndex(-1) {}
};

struct MosModel {
    std::string name;
    bool isP = false;
    double VT = 0.7, MU = 1e-3, COX = 1e-3, LAMBDA = 0.0, CJO = 0.0;
};

class Circuit {
public:
    std::vector<Node> nodes;
    std::vector<std::shared_ptr<Element>> elements;
    std::unordered_map<std::string, int> nodeNameToId;
    std::unordered_map<std::string, MosModel> mosModels;

    int getOrCreateNode(const std::string& name);
    int numNodeEquations() const;
    int numVoltageBranches() const;
    int numUnknowns() const;
    void assignEquationIndices();

    void addResistor(const std::string& name, const std::string& n1, const std::string& n2, double value);
    void addCapacitor(const std::string& name, const std::string& n1, const std::string& n2, double value);
    void addInductor(const std::string& name, const std::string& n1, const std::string& n2, double value);
    void addCurrentSource(const std::string& name, const std::string& np, const std::string& nm,
                          const SourceSpec& spec);
    void addVoltageSource(const std::string& name, const std::string& np, const std::string& nm,
                          const SourceSpec& spec);
    void addMosfet(const std::string& name, const std::string& nd, const std::string& ng, const std::string& ns,
                   const std::string& modelId, double W, double L);
    void addMosModel(const MosModel& m);
    const MosModel* findMosModel(const std::string& id) const;
    void printConnectivity() const;

private:
    void attach(const std::shared_ptr<Element>& e, std::initializer_list<int> nodeIds);
};
