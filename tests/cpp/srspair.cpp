// The srspair writer of the C++ host mirror (include/swb200.hpp) on aligned strings read from a TSV file:
//   kind \t id \t start1 \t start2 \t score \t aln1 \t aln2      -> prints each body preceded by "@@ kind id\n".
// No device work: only the built-in matrix is fetched from the library.
#include <fstream>
#include <iostream>
#include <sstream>

#include "swb200.hpp"

int main(int argc, char **argv)
{
    if (argc < 2) { std::cerr << "usage: srspair cases.tsv\n"; return 2; }
    const swb200::SubstitutionMatrix m = swb200::SubstitutionMatrix::get_blosum62_matrix();
    std::ifstream f(argv[1]);
    std::string line;
    while (std::getline(f, line)) {
        std::vector<std::string> c;
        std::stringstream ss(line);
        std::string cell;
        while (std::getline(ss, cell, '\t')) c.push_back(cell);
        if (c.size() < 7) continue;
        std::cout << "@@ " << c[0] << " " << c[1] << "\n"
                  << swb200::srspair_body(c[5], c[6], std::stof(c[4]), m, 10.0f, 0.5f, std::stoll(c[2]), std::stoll(c[3]));
    }
    return 0;
}
