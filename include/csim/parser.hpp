// parser.hpp — SPICE-like .sp front end.  Entry points and behaviour of the reference's
// include/parser.hpp / src/parser.cpp: comments (* ; full line, $ inline), '+' continuation,
// .MODEL cards first, devices R C L V I M, cards .op .dc .tran .ac .print .model .hb .plotnv
// .plotnc; statement errors go to std::cerr and the statement is skipped; parseStream
// still returns true (src/parser.cpp:51-57).
#pragma once
#include <iosfwd>
#include <string>
#include <vector>

#include "circuit.hpp"
#include "sim.hpp"

class NetlistParser {
public:
    NetlistParser(Circuit& circuit, SimulationConfig& simConfig);
    bool parseFile(const std::string& filename);
    bool parseStream(std::istream& in, const std::string& originName = "<stream>");

private:
    struct Statement {
        int lineNo = 0;
        std::string raw;
        std::vector<std::string> tokens;
    };
    Circuit& ckt;
    SimulationConfig& sim;
    std::string sourceName;
    std::vector<Statement> stmts;

    void lex(std::istream& in);
    void parseStatements();
    void device(const Statement& st);
    void card(const Statement& st);
    void twoTerminal(const Statement& st, char kind);
    void voltageSource(const Statement& st);
    void currentSource(const Statement& st);
    void mosfet(const Statement& st);
    void model(const Statement& st);
    void plotCard(const Statement& st, bool current);
    void printCard(const Statement& st);
    static ProbeSpec probeFromToken(const std::string& token);
    bool complain(const Statement& st, const std::string& what) const;
};

inline bool parseNetlist(const std::string& filename, Circuit& ckt, SimulationConfig& sim) {
    NetlistParser parser(ckt, sim);
    bool ok = parser.parseFile(filename);
    sim.ensureDefaultOp();
    return ok;
}
