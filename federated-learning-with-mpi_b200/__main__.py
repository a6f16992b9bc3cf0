from fedb200.cli import main

if __name__ == "__main__":
    raise SystemExit(main())
