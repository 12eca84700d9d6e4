ssh test@127.0.0.1 "kill -9 9177"
ssh test@127.0.0.1 "kill -9 9178"
kill -9 8873
