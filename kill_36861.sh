ssh test@127.0.0.1 "kill -9 8899"
ssh test@127.0.0.1 "kill -9 8900"
kill -9 8873
