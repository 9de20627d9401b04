// element.hpp — circuit elements.  Class names, constructors and the stamp() signature are
// the reference's (include/element.hpp); the new const getters are the superset SURVEY.md
// §8(b) asks for (the reference keeps R, Vth, K, lambda, isP inaccessible).
//
// stamp() keeps the reference's meaning — ADD this element's contribution for iterate x
// into G and I — but the numbers come from the device: the element is lowered to a
// one-element csim_topology and assembled by the CUDA stamp kernel (no CPU model code).
// The batched analyses never call it; they lower the whole Circuit once (lower.hpp).
#pragma once
#include <memory>
#include <string>
#include <vector>

#include "dense.hpp"
#include "sim.hpp"

class Circuit;

class Element {
protected:
    std::string name;
    std::vector<int> nodeIds;  // indices into Circuit::nodes

public:
    Element(const std::string& n, const std::vector<int>& nodes) : name(n), nodeIds(nodes) {}
    virtual ~Element() {}

    const std::string& getName() const { return name; }
    const std::vector<int>& getNodeIds() const { return nodeIds; }

    // csim_elem_type of this element and its parameter doubles in C-ABI order
    virtual int abiType() const = 0;
    virtual int abiFlags() const { return 0; }
    virtual int abiBranch() const { return -1; }
    virtual void abiParams(std::vector<double>& out) const = 0;

    // reference: include/element.hpp:28-31.  Throws std::runtime_error when no CUDA device.
    virtual void stamp(Eigen::MatrixXd& G, Eigen::VectorXd& I, const Circuit& ckt, const Eigen::VectorXd& x,
                       const AnalysisContext& ctx) const;

    // AC stamps are outside the Newton/MNA path (no reference driver calls them): no-op.
    virtual void stampAC(Eigen::MatrixXcd&, Eigen::VectorXcd&, const Circuit&, double) const {}
};

class Resistor : public Element {
    double R;

public:
    Resistor(const std::string& n, int n1, int n2, double r) : Element(n, {n1, n2}), R(r) {}
    double getR() const { return R; }
    int abiType() const override;
    void abiParams(std::vector<double>& out) const override { out.push_back(R); }
};

class CurrentSource : public Element {
    SourceSpec spec;

public:
    CurrentSource(const std::string& n, int np, int nm, const SourceSpec& s) : Element(n, {np, nm}), spec(s) {}
    const SourceSpec& setSpec() const { return spec; }  // (sic) the reference's name for this getter
    const SourceSpec& getSpec() const { return spec; }
    void replaceSpec(const SourceSpec& s) { spec = s; }  // superset: PULSE / PWL sources (the parser only makes SIN)
    int abiType() const override;
    int abiFlags() const override;
    void abiParams(std::vector<double>& out) const override;
};

class VoltageSource : public Element {
    SourceSpec spec;
    int branchEqIndex;

public:
    VoltageSource(const std::string& n, int np, int nm, const SourceSpec& s)
        : Element(n, {np, nm}), spec(s), branchEqIndex(-1) {}
    void setBranchEqIndex(int idx) { branchEqIndex = idx; }
    int getBranchEqIndex() const { return branchEqIndex; }
    const SourceSpec& getSpec() const { return spec; }
    void replaceSpec(const SourceSpec& s) { spec = s; }  // superset: PULSE / PWL sources (the parser only makes SIN)
    int abiType() const override;
    int abiFlags() const override;
    int abiBranch() const override { return branchEqIndex; }
    void abiParams(std::vector<double>& out) const override;
};

class CapacitorElement : public Element {
    double C;

public:
    CapacitorElement(const std::string& n, int n1, int n2, double c) : Element(n, {n1, n2}), C(c) {}
    double getC() const { return C; }
    int abiType() const override;
    void abiParams(std::vector<double>& out) const override { out.push_back(C); }
};

class Inductor : public Element {
    double L;
    int branchEqIndex;

public:
    Inductor(const std::string& n, int n1, int n2, double l) : Element(n, {n1, n2}), L(l), branchEqIndex(-1) {}
    void setBranchEqIndex(int idx) { branchEqIndex = idx; }
    int getBranchEqIndex() const { return branchEqIndex; }
    double getL() const { return L; }
    int abiType() const override;
    int abiBranch() const override { return branchEqIndex; }
    void abiParams(std::vector<double>& out) const override { out.push_back(L); }
};

class MosfetBase : public Element {
protected:
    bool isP;
    double Vth, K, lambda, Cj0;

public:
    MosfetBase(const std::string& n, int nd, int ng, int ns, int nb, bool isPchannel, double Vth_, double K_,
               double lambda_, double Cj0_)
        : Element(n, {nd, ng, ns, nb}), isP(isPchannel), Vth(Vth_), K(K_), lambda(lambda_), Cj0(Cj0_) {}
    double getCj0() const { return Cj0; }
    double getVth() const { return Vth; }
    double getK() const { return K; }
    double getLambda() const { return lambda; }
    bool isPChannel() const { return isP; }
    int abiType() const override;
    int abiFlags() const override;
    void abiParams(std::vector<double>& out) const override {
        out.insert(out.end(), {Vth, K, lambda, Cj0});
    }
};

class NMosElement : public MosfetBase {
public:
    NMosElement(const std::string& n, int nd, int ng, int ns, int nb, double Vth_, double K_, double lambda_,
                double Cj0_)
        : MosfetBase(n, nd, ng, ns, nb, false, Vth_, K_, lambda_, Cj0_) {}
};

class PMosElement : public MosfetBase {
public:
    PMosElement(const std::string& n, int nd, int ng, int ns, int nb, double Vth_, double K_, double lambda_,
                double Cj0_)
        : MosfetBase(n, nd, ng, ns, nb, true, Vth_, K_, lambda_, Cj0_) {}
};
