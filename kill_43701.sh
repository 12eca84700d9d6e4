ssh test@127.0.0.1 "kill -9 9255"
ssh test@127.0.0.1 "kill -9 9256"
kill -9 8873
