// TEST INFRASTRUCTURE — not part of the product path.
//
// Trimmed driver for the UNMODIFIED reference GMM path: the `train` and `recognize` branches of
// /root/reference/src/sietill/SieTill.cpp:29-50,80-135 without the neural-network includes (cblas.h
// is absent from this image, so NeuralNetwork/NNTraining cannot be compiled).  Built by oracle/Makefile
// into oracle/_ref/sietill_ref; used in this container to run the reference end to end
// (config 1 of BASELINE.json) and to generate golden fixtures.
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>

#include "Config.hpp"
#include "Corpus.hpp"
#include "Recognizer.hpp"
#include "SignalAnalysis.hpp"
#include "Training.hpp"

int main(int argc, const char* argv[]) {
  if (argc < 2) {
    std::cerr << "usage: " << argv[0] << " <config.json> [action]" << std::endl;
    return EXIT_FAILURE;
  }
  const Configuration config(argv[1]);
  std::string action = ParameterString("action", "")(config);
  if (argc >= 3) action = argv[2];
  const std::string feature_path = ParameterString("feature-path", "")(config);
  const std::string norm_path    = ParameterString("normalization-path", "")(config);
  const bool max_approx          = ParameterBool("max-approx", true)(config);
  const std::string pooling      = ParameterString("pooling", "")(config);

  Lexicon lexicon = build_sietill_lexicon();
  CorpusDescription description(config);
  description.read(lexicon);
  SignalAnalysis analyzer(config);
  if (!norm_path.empty()) {
    std::ifstream in(norm_path.c_str(), std::ios_base::in);
    if (!in.good()) { std::cerr << "cannot open normalization file" << std::endl; return EXIT_FAILURE; }
    analyzer.read_normalization_file(in);
  }
  MixtureModel::VarianceModel vm = MixtureModel::NO_POOLING;
  if (pooling == "mixture") vm = MixtureModel::MIXTURE_POOLING;
  else if (pooling == "global") vm = MixtureModel::GLOBAL_POOLING;
  else if (pooling != "none") { std::cerr << "bad pooling: " << pooling << std::endl; return EXIT_FAILURE; }

  Corpus corpus;
  corpus.read(description, feature_path, analyzer);
  TdpModel tdp(config, lexicon.get_silence_automaton()[0ul]);
  MixtureModel mixtures(config, analyzer.n_features_total, lexicon.num_states(), vm, max_approx);
  if (action == "train") {
    Trainer trainer(config, lexicon, mixtures, tdp, max_approx);
    trainer.train(corpus);
  } else if (action == "recognize") {
    Recognizer recognizer(config, lexicon, mixtures, tdp);
    recognizer.recognize(corpus);
  } else {
    std::cerr << "unsupported action: " << action << std::endl;
    return EXIT_FAILURE;
  }
  return EXIT_SUCCESS;
}
