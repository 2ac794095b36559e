"""Special-token constants of the reference tokenizer (reference perceiver/tokenizer.py:10-19).

Only the ids matter to the hot path (TextMasking); WordPiece training / loading is host-side text
preprocessing and stays with the HuggingFace `tokenizers` package, imported lazily."""

PAD_TOKEN, PAD_TOKEN_ID = '[PAD]', 0
UNK_TOKEN, UNK_TOKEN_ID = '[UNK]', 1
MASK_TOKEN, MASK_TOKEN_ID = '[MASK]', 2
SPECIAL_TOKENS = [PAD_TOKEN, UNK_TOKEN, MASK_TOKEN]


def load_tokenizer(path):
    from tokenizers import Tokenizer
    return Tokenizer.from_file(path)


def save_tokenizer(tokenizer, path):
    tokenizer.save(path)


def create_tokenizer(*normalizer):
    from tokenizers import Tokenizer, decoders
    from tokenizers.models import WordPiece
    from tokenizers.normalizers import NFD, Lowercase, Sequence, StripAccents
    from tokenizers.pre_tokenizers import Whitespace
    tok = Tokenizer(WordPiece(unk_token=UNK_TOKEN))
    tok.normalizer = Sequence([*normalizer, NFD(), Lowercase(), StripAccents()])
    tok.pre_tokenizer = Whitespace()
    tok.decoder = decoders.WordPiece()
    return tok


def train_tokenizer(tokenizer, data, vocab_size):
    from tokenizers.trainers import WordPieceTrainer
    tokenizer.train_from_iterator(data, WordPieceTrainer(vocab_size=vocab_size, special_tokens=SPECIAL_TOKENS))
