ssh test@127.0.0.1 "kill -9 9127"
ssh test@127.0.0.1 "kill -9 9128"
ssh test@127.0.0.1 "kill -9 9129"
ssh test@127.0.0.1 "kill -9 9130"
kill -9 8873
