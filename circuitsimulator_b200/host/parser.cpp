// parser.cpp — .sp netlist reader (behavioural contract: reference src/parser.cpp).
#include "csim/parser.hpp"

#include <cctype>
#include <fstream>
#include <iostream>
#include <sstream>
#include <stdexcept>

namespace {

std::string trim(const std::string& s) {
    const char* ws = " \t\r\n";
    const auto a = s.find_first_not_of(ws);
    if (a == std::string::npos) return "";
    const auto b = s.find_last_not_of(ws);
    return s.substr(a, b - a + 1);
}
std::string dropDollarComment(const std::string& s) {
    const auto p = s.find('$');
    return p == std::string::npos ? s : s.substr(0, p);
}
bool keyword(const std::string& tok, const char* kw) { return toLower(tok) == kw; }

// first '(' and last ')' of a probe token
void parenSpan(const std::string& s, int& l, int& r) {
    l = r = -1;
    for (int i = 0; i < static_cast<int>(s.size()); ++i) {
        if (s[i] == '(' && l < 0) l = i;
        if (s[i] == ')') r = i;
    }
}

}  // namespace

NetlistParser::NetlistParser(Circuit& circuit, SimulationConfig& simConfig) : ckt(circuit), sim(simConfig) {}

bool NetlistParser::parseFile(const std::string& filename) {
    std::ifstream in(filename);
    if (!in) {
        std::cerr << "cannot open netlist file " << filename << "\n";
        return false;
    }
    return parseStream(in, filename);
}

bool NetlistParser::parseStream(std::istream& in, const std::string& originName) {
    sourceName = originName;
    lex(in);
    parseStatements();
    return true;  // statement-level errors are reported and skipped, never fatal
}

bool NetlistParser::complain(const Statement& st, const std::string& what) const {
    std::cerr << "Line " << st.lineNo << ": " << what << ": " << st.raw << "\n";
    return false;
}

// Physical lines -> logical statements: '$' starts an inline comment, a line whose first
// character is '*' or ';' is dropped, a leading '+' glues the rest onto the open statement.
void NetlistParser::lex(std::istream& in) {
    stmts.clear();
    std::string open;
    int openLine = 0, lineNo = 0;
    auto emit = [&]() {
        const std::string text = trim(dropDollarComment(open));
        open.clear();
        if (text.empty()) return;
        Statement st;
        st.lineNo = openLine;
        st.raw = text;
        std::istringstream words(text);
        for (std::string w; words >> w;) st.tokens.push_back(w);
        if (!st.tokens.empty()) stmts.push_back(std::move(st));
    };
    for (std::string line; std::getline(in, line);) {
        ++lineNo;
        const std::string text = trim(dropDollarComment(line));
        if (text.empty() || text[0] == '*' || text[0] == ';') continue;
        if (text[0] == '+') {
            const std::string rest = ltrim(text.substr(1));
            if (open.empty()) {
                openLine = lineNo;
                open = rest;
            } else {
                open += " " + rest;
            }
            continue;
        }
        if (!open.empty()) emit();
        openLine = lineNo;
        open = text;
    }
    if (!open.empty()) emit();
}

void NetlistParser::parseStatements() {
    for (const Statement& st : stmts)  // models first so that M cards can appear anywhere
        if (keyword(st.tokens[0], ".model")) model(st);
    for (const Statement& st : stmts) {
        if (st.tokens[0][0] == '.') {
            if (!keyword(st.tokens[0], ".model")) card(st);
        } else {
            device(st);  // a title line ends up here too and is reported as unsupported, like the reference
        }
    }
    sim.ensureDefaultOp();
}

void NetlistParser::device(const Statement& st) {
    switch (std::toupper(static_cast<unsigned char>(st.tokens[0][0]))) {
        case 'R': twoTerminal(st, 'R'); break;
        case 'C': twoTerminal(st, 'C'); break;
        case 'L': twoTerminal(st, 'L'); break;
        case 'V': voltageSource(st); break;
        case 'I': currentSource(st); break;
        case 'M': mosfet(st); break;
        default: complain(st, "unsupported element or syntax");
    }
}

void NetlistParser::twoTerminal(const Statement& st, char kind) {
    const auto& t = st.tokens;
    const char* what = kind == 'R' ? "resistor" : kind == 'C' ? "capacitor" : "inductor";
    if (t.size() < 4) {
        complain(st, std::string("invalid ") + what);
        return;
    }
    double value;
    try {
        value = parseSpiceNumber(t[3]);
    } catch (const std::exception& e) {
        complain(st, std::string("cannot parse ") + kind + " value (" + e.what() + ")");
        return;
    }
    if (kind == 'R') ckt.addResistor(t[0], t[1], t[2], value);
    else if (kind == 'C') ckt.addCapacitor(t[0], t[1], t[2], value);
    else ckt.addInductor(t[0], t[1], t[2], value);
}

// Vname n+ n- <value> | DC <value> | SIN v0 va freq [td [phi]]   (value forms may be followed by SIN ...)
// NOTE the 5th SIN number is a DELAY in seconds and phi is in radians (reference quirk).
void NetlistParser::voltageSource(const Statement& st) {
    const auto& t = st.tokens;
    if (t.size() < 4) {
        complain(st, "invalid voltage source");
        return;
    }
    SourceSpec spec;
    std::size_t next;
    try {
        if (t.size() >= 5 && keyword(t[3], "dc")) {
            spec.dcValue = parseSpiceNumber(t[4]);
            next = 5;
        } else if (keyword(t[3], "sin")) {
            next = 3;
        } else {
            spec.dcValue = parseSpiceNumber(t[3]);
            next = 4;
        }
    } catch (const std::exception& e) {
        complain(st, std::string("cannot parse V DC value (") + e.what() + ")");
        return;
    }
    if (next < t.size() && keyword(t[next], "sin")) {
        if (t.size() < next + 4) {
            complain(st, "SIN needs at least 3 parameters (v0 va freq)");
        } else {
            try {
                SinSpec s;
                s.v0 = parseSpiceNumber(t[next + 1]);
                s.va = parseSpiceNumber(t[next + 2]);
                s.freq = parseSpiceNumber(t[next + 3]);
                if (t.size() > next + 4) s.td = parseSpiceNumber(t[next + 4]);
                if (t.size() > next + 5) s.phi = parseSpiceNumber(t[next + 5]);
                spec.tran.type = WaveformType::SIN;
                spec.tran.sine = s;
            } catch (const std::exception& e) {
                complain(st, std::string("cannot parse SIN parameters (") + e.what() + ")");
            }
        }
    }
    ckt.addVoltageSource(t[0], t[1], t[2], spec);
}

void NetlistParser::currentSource(const Statement& st) {
    const auto& t = st.tokens;
    if (t.size() < 4) {
        complain(st, "invalid current source");
        return;
    }
    SourceSpec spec;
    try {
        spec.dcValue = parseSpiceNumber((t.size() >= 5 && keyword(t[3], "dc")) ? t[4] : t[3]);
    } catch (const std::exception& e) {
        complain(st, std::string("cannot parse I value (") + e.what() + ")");
        return;
    }
    ckt.addCurrentSource(t[0], t[1], t[2], spec);
}

// "M name nd ng ns model W L" (7 tokens) or "M name nd ng ns p|n W L modelId" (8 tokens; the
// p/n letter is ignored, the channel type comes from the model's VT sign).
void NetlistParser::mosfet(const Statement& st) {
    const auto& t = st.tokens;
    if (t.size() != 7 && t.size() != 8) {
        complain(st, "invalid MOSFET");
        return;
    }
    const std::string modelId = t.size() == 7 ? t[4] : t.back();
    double W, L;
    try {
        W = parseSpiceNumber(t[5]);
        L = parseSpiceNumber(t[6]);
    } catch (const std::exception& e) {
        complain(st, std::string("cannot parse MOS W/L (") + e.what() + ")");
        return;
    }
    ckt.addMosfet(t[0], t[1], t[2], t[3], modelId, W, L);
}

void NetlistParser::model(const Statement& st) {
    const auto& t = st.tokens;
    if (t.size() < 4) {
        complain(st, "invalid .MODEL");
        return;
    }
    MosModel m;
    m.name = t[1];
    for (std::size_t i = 2; i + 1 < t.size(); i += 2) {
        double v;
        try {
            v = parseSpiceNumber(t[i + 1]);
        } catch (const std::exception& e) {
            complain(st, "cannot parse .MODEL param " + t[i] + " (" + e.what() + ")");
            return;
        }
        const std::string key = toLower(t[i]);
        if (key == "vt") m.VT = v;
        else if (key == "mu") m.MU = v;
        else if (key == "cox") m.COX = v;
        else if (key == "lambda") m.LAMBDA = v;
        else if (key == "cj0" || key == "cjo") m.CJO = v;
    }
    if (m.VT < 0.0) {  // the sign of VT selects PMOS; the magnitude is stored
        m.isP = true;
        m.VT = -m.VT;
    }
    ckt.addMosModel(m);
}

// ---- control cards -----------------------------------------------------------------
void NetlistParser::card(const Statement& st) {
    const auto& t = st.tokens;
    const std::string head = toLower(t[0]);
    try {
        if (head == ".op") {
            sim.doOp = true;
        } else if (head == ".dc") {
            if (t.size() < 5) {
                complain(st, "invalid .DC syntax");
                return;
            }
            DCSweepConfig dc;
            dc.sourceName = t[1];
            dc.start = parseSpiceNumber(t[2]);
            dc.stop = parseSpiceNumber(t[3]);
            dc.step = parseSpiceNumber(t[4]);
            sim.dcSweeps.push_back(dc);
        } else if (head == ".tran") {
            if (t.size() < 3) {
                complain(st, "invalid .TRAN syntax");
                return;
            }
            TranConfig cfg;
            cfg.tstep = parseSpiceNumber(t[1]);
            cfg.tstop = parseSpiceNumber(t[2]);
            cfg.tstart = t.size() >= 4 ? parseSpiceNumber(t[3]) : 0.0;
            cfg.enabled = true;
            sim.tran = cfg;
        } else if (head == ".ac") {
            if (t.size() < 5) {
                complain(st, "invalid .AC syntax");
                return;
            }
            AcConfig cfg;
            const std::string sw = toLower(t[1]);
            cfg.sweepType = sw == "lin" ? AcSweepType::LIN : sw == "oct" ? AcSweepType::OCT : AcSweepType::DEC;
            cfg.nPoints = std::stoi(t[2]);
            cfg.fstart = parseSpiceNumber(t[3]);
            cfg.fstop = parseSpiceNumber(t[4]);
            cfg.enabled = true;
            sim.ac = cfg;
        } else if (head == ".hb") {
            if (t.size() < 3) {
                complain(st, "invalid .hb syntax");
                return;
            }
            HbConfig cfg;
            cfg.f0 = parseSpiceNumber(t[1]);
            cfg.nHarm = std::stoi(t[2]);
            cfg.enabled = true;
            sim.hb = cfg;
        } else if (head == ".mc") {  // superset card (montecarlo.hpp); the reference reports it as unsupported
            if (t.size() < 2) {
                complain(st, "invalid .MC syntax");
                return;
            }
            MonteCarloConfig cfg;
            cfg.instances = std::stoll(t[1]);
            if (t.size() >= 3) cfg.seed = std::stoull(t[2]);
            if (t.size() >= 4) cfg.rel = parseSpiceNumber(t[3]);
            if (cfg.instances <= 0 || !(cfg.rel >= 0.0)) {
                complain(st, "invalid .MC syntax");
                return;
            }
            cfg.enabled = true;
            sim.mc = cfg;
        } else if (head == ".print") {
            printCard(st);
        } else if (head == ".plotnv") {
            plotCard(st, false);
        } else if (head == ".plotnc") {
            plotCard(st, true);
        } else {
            complain(st, "unsupported control card");
        }
    } catch (const std::exception& e) {
        complain(st, std::string("cannot parse ") + head + " arguments (" + e.what() + ")");
    }
}

ProbeSpec NetlistParser::probeFromToken(const std::string& token) {
    ProbeSpec p;
    p.expr = token;
    if (token.empty()) return p;
    int l, r;
    parenSpan(token, l, r);
    const bool hasInside = l >= 0 && r > l + 1;
    const std::string inside = hasInside ? token.substr(l + 1, r - l - 1) : std::string();
    const int c0 = std::toupper(static_cast<unsigned char>(token[0]));
    if (c0 == 'V') {
        p.kind = ProbeKind::NodeVoltage;
        if (hasInside) {
            const auto comma = inside.find(',');
            if (comma == std::string::npos) {
                p.node1 = trim(inside);
            } else {
                p.node1 = trim(inside.substr(0, comma));
                p.node2 = trim(inside.substr(comma + 1));
                p.kind = ProbeKind::DiffVoltage;
            }
        }
    } else if (c0 == 'I') {
        p.kind = ProbeKind::BranchCurrent;
        if (hasInside) p.eleName = trim(inside);
    }
    return p;
}

void NetlistParser::printCard(const Statement& st) {
    const auto& t = st.tokens;
    if (t.size() < 3) {
        complain(st, "invalid .PRINT");
        return;
    }
    static const struct {
        const char* word;
        AnalysisType type;
    } kinds[] = {{"op", AnalysisType::OP}, {"dc", AnalysisType::DC}, {"ac", AnalysisType::AC},
                 {"tran", AnalysisType::TRAN}, {"hb", AnalysisType::HB}};
    PrintCommand pc;
    const std::string which = toLower(t[1]);
    for (const auto& k : kinds)
        if (which == k.word) pc.analysis = k.type;
    if (pc.analysis == AnalysisType::NONE) {
        complain(st, "unknown analysis type in .PRINT");
        return;
    }
    for (std::size_t i = 2; i < t.size(); ++i) pc.probes.push_back(probeFromToken(t[i]));
    sim.printCommands.push_back(std::move(pc));
}

// .PLOTNV node...   -> node-voltage probes ;  .PLOTNC elem[(port)]... -> branch-current probes.
// Not bound to an analysis type (AnalysisType::NONE).
void NetlistParser::plotCard(const Statement& st, bool current) {
    const auto& t = st.tokens;
    if (t.size() < 2) {
        complain(st, current ? "invalid .PLOTNC" : "invalid .PLOTNV");
        return;
    }
    PrintCommand pc;
    for (std::size_t i = 1; i < t.size(); ++i) {
        if (t[i].empty()) continue;
        if (!current) {
            pc.probes.push_back(probeFromToken("V(" + t[i] + ")"));
            continue;
        }
        ProbeSpec p;
        p.kind = ProbeKind::BranchCurrent;
        p.expr = t[i];
        int l, r;
        parenSpan(t[i], l, r);
        if (l < 0) {
            p.eleName = t[i];
        } else {
            p.eleName = trim(t[i].substr(0, l));
            p.elePort = r > l ? trim(t[i].substr(l + 1, r - l - 1)) : std::string();
        }
        pc.probes.push_back(std::move(p));
    }
    if (!pc.probes.empty()) sim.printCommands.push_back(std::move(pc));
}
